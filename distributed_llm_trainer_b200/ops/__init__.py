"""CUDA kernels (ctypes bindings in ``kernels``), the fused engine, the flat AdamW and the PyTorch
oracles (``reference``).  Importing this package never touches the GPU or the native library."""
from . import _lib

__all__ = ["_lib"]
