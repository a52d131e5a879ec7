"""Checkpoint file IO: atomic writes, tolerant loads, latest-checkpoint discovery.

Layout parity: top-level keys ``model, optimizer, global_step, tokens_seen,
model_config, training_config`` (+ ``fsdp_config`` for FSDP) -- reference ddp_trainer.py:408-415 and
fsdp_trainer.py:453-461.  Differences, on purpose:

* configs are stored as plain ``dict`` (the reference pickles ``__main__.TrainingConfig`` /
  ``__main__.FSDPConfig``, which makes its FSDP checkpoints unloadable by its own ``infer.py`` --
  SURVEY D9); ``load_checkpoint_file`` still reads the reference's pickled-class files by mapping
  unknown classes to an attribute bag;
* writes go to ``path.tmp`` then ``os.replace`` (the reference writes in place, no barrier);
* RNG / dropout counter state is saved under ``extra`` so a resumed run continues the same streams.
"""
from __future__ import annotations

import glob
import io
import os
import pickle
import re
from typing import Any, Dict, Optional

import torch


class _AttrBag:
    """Stand-in for classes that cannot be imported when reading foreign checkpoints."""

    def __init__(self, *a, **k):
        pass

    def __setstate__(self, state):
        if isinstance(state, dict):
            self.__dict__.update(state)


class _TolerantUnpickler(pickle.Unpickler):
    def find_class(self, module: str, name: str):
        if module in ("models.config", "src.models.config") and name == "GPTConfig":
            from ..models.config import GPTConfig

            return GPTConfig
        try:
            return super().find_class(module, name)
        except Exception:
            return type(name, (_AttrBag,), {"__module__": module})


class _TolerantPickle:
    """``pickle_module`` shim for ``torch.load``."""
    __name__ = "tolerant_pickle"
    Unpickler = _TolerantUnpickler
    load = staticmethod(lambda f, **kw: _TolerantUnpickler(f, **kw).load())
    loads = staticmethod(lambda b, **kw: _TolerantUnpickler(io.BytesIO(b), **kw).load())
    dump = staticmethod(pickle.dump)
    dumps = staticmethod(pickle.dumps)
    PickleError = pickle.PickleError
    UnpicklingError = pickle.UnpicklingError


def atomic_save(obj: Dict[str, Any], path: str) -> None:
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)
    tmp = f"{path}.tmp.{os.getpid()}"
    torch.save(obj, tmp)
    os.replace(tmp, path)


def load_checkpoint_file(path: str, map_location: Any = "cpu") -> Dict[str, Any]:
    """Read our checkpoints and the reference's (pickled config classes included)."""
    try:
        return torch.load(path, map_location=map_location, weights_only=False)
    except Exception:
        return torch.load(path, map_location=map_location, weights_only=False, pickle_module=_TolerantPickle)


def config_to_plain(cfg: Any) -> Any:
    if cfg is None or isinstance(cfg, dict):
        return cfg
    if hasattr(cfg, "to_dict"):
        return cfg.to_dict()
    return dict(getattr(cfg, "__dict__", {}))


def find_latest_checkpoint(directory: str) -> Optional[str]:
    """``final.pt`` if present, else the highest ``step_N.pt``."""
    final = os.path.join(directory, "final.pt")
    if os.path.exists(final):
        return final
    best, best_step = None, -1
    for p in glob.glob(os.path.join(directory, "step_*.pt")):
        m = re.search(r"step_(\d+)\.pt$", p)
        if m and int(m.group(1)) > best_step:
            best, best_step = p, int(m.group(1))
    return best
