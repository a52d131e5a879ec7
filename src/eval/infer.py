#!/usr/bin/env python3
"""Same command line as the reference's src/eval/infer.py; the implementation lives in
distributed_llm_trainer_b200.eval.infer."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from distributed_llm_trainer_b200.eval.infer import *  # noqa: F401,F403,E402
from distributed_llm_trainer_b200.eval.infer import main  # noqa: E402

if __name__ == "__main__":
    main()
