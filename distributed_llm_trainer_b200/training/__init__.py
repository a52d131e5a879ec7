from .config import FSDPConfig, TrainingConfig, load_yaml_config
from .lr import cosine_lr

__all__ = ["FSDPConfig", "TrainingConfig", "load_yaml_config", "cosine_lr"]
