"""Import-path shim: reference checkpoints pickle ``models.config.GPTConfig``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from distributed_llm_trainer_b200.models.config import GPTConfig  # noqa: F401,E402
