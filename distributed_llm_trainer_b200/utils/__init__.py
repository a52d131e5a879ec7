from .checkpoint import atomic_save, config_to_plain, find_latest_checkpoint, load_checkpoint_file
from .dist import DistInfo, init_distributed
from .logging import MetricsLogger, format_ddp_line, format_fsdp_line

__all__ = [n for n in dir() if not n.startswith("_")]
