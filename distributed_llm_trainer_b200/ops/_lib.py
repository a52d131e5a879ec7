"""Loader (and in-tree builder) for ``libb200trainer.so``.

The native library is plain ``nvcc``-compiled C ABI (``include/b200_trainer.h``) loaded with
``ctypes``: no torch headers are involved, so a full rebuild takes seconds and the same ``.so``
serves Python, C and C++ callers.  It is built **in-tree** (``ops/libb200trainer.so``) so the
artefact travels with a snapshot of the repository.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import threading
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
_REPO = os.path.dirname(os.path.dirname(_HERE))
LIB_PATH = os.path.join(_HERE, "libb200trainer.so")
CSRC = os.path.join(_REPO, "csrc")

_lib: Optional[ctypes.CDLL] = None
_lock = threading.Lock()


class NativeLibraryError(RuntimeError):
    pass


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every ``csrc/*.cu`` for sm_100a (``-gencode arch=compute_100a,code=sm_100a -lineinfo``)."""
    if force and os.path.exists(LIB_PATH):
        os.remove(LIB_PATH)
    jobs = str(max(1, min(8, os.cpu_count() or 1)))
    proc = subprocess.run(["make", "-C", CSRC, "-j", jobs], capture_output=True, text=True)
    if verbose or proc.returncode != 0:
        print(proc.stdout)
        print(proc.stderr)
    if proc.returncode != 0 or not os.path.exists(LIB_PATH):
        raise NativeLibraryError(f"building {LIB_PATH} failed:\n{proc.stdout[-2000:]}\n{proc.stderr[-2000:]}")
    return LIB_PATH


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> ctypes.CDLL:
    """The loaded library.  Raises (never falls back) when it is missing: on a GPU box a silent
    PyTorch fallback would make every benchmark and test meaningless."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise NativeLibraryError(
                        f"{LIB_PATH} not found -- run `python -c 'import __graft_entry__ as g; g.build()'` "
                        f"or `make -C csrc` first")
                handle = ctypes.CDLL(LIB_PATH)
                handle.b200_last_error.restype = ctypes.c_char_p
                _lib = handle
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().b200_last_error()
        raise NativeLibraryError(f"{what or 'b200 call'} failed ({rc}): {msg.decode() if msg else '?'}")
