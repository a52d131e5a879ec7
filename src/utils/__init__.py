from distributed_llm_trainer_b200.utils import *  # noqa: F401,F403
