"""distributed_llm_trainer_b200 -- a Blackwell (sm_100a) native GPT training framework.

Layers:  models/  ops/ (CUDA kernels + engine)  parallel/ (DDP + FSDP engines,
NVLink comm runtime)  training/ (trainers, CLI, config)  data/  eval/  utils/.
"""
__version__ = "0.1.0"
