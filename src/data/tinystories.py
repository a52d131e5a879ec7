"""Import-path shim for ``from data.tinystories import create_tinystories_dataloader``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from distributed_llm_trainer_b200.data.text import (TinyStoriesConfig, TinyStoriesDataset,  # noqa: F401,E402
                                                     TinyStoriesIterableDataset, create_tinystories_dataloader)
