"""Import-path shim for ``from data.openwebtext import create_openwebtext_dataloader``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from distributed_llm_trainer_b200.data.text import (OpenWebTextConfig, OpenWebTextDataset,  # noqa: F401,E402
                                                     OpenWebTextIterableDataset, create_openwebtext_dataloader)
