"""Import-path shim for ``from models.gpt import GPT, count_parameters, TransformerBlock``."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from distributed_llm_trainer_b200.models.gpt import *  # noqa: F401,F403,E402
from distributed_llm_trainer_b200.models.gpt import GPT, TransformerBlock, count_parameters  # noqa: F401,E402
