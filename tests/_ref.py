"""Locate the unmodified reference sources (oracle for parity tests)."""
import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_CANDIDATES = [os.path.join(ROOT, "baseline", "_ref"), "/root/reference/src"]


def ref_path():
    for c in _CANDIDATES:
        if os.path.isdir(os.path.join(c, "models")):
            return c
    return None


def import_ref(module: str):
    """Import a reference module (``models.gpt``, ``training.ddp_trainer`` ...) or skip the test."""
    p = ref_path()
    if p is None:
        pytest.skip("reference sources not available")
    if p not in sys.path:
        sys.path.insert(0, p)
    for name in list(sys.modules):
        root = name.split(".")[0]
        if root in ("models", "data", "training") and not (getattr(sys.modules[name], "__file__", "") or "").startswith(p):
            del sys.modules[name]
    return importlib.import_module(module)
