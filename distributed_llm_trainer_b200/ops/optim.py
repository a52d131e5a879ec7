"""AdamW over the flat parameter store.

Math is ``torch.optim.AdamW`` (decoupled decay, bias-corrected, eps outside the sqrt) as used by
the reference (ddp_trainer.py:229-234 fused, fsdp_trainer.py:338-343 foreach).  One step touches
a contiguous ``[lo, hi)`` range of the flat space -- the whole buffer on one GPU, this rank's
slice under the sharded DDP / FSDP engines -- in a single kernel launch that also

* applies the global-norm clip coefficient read from device memory (no host sync),
* rewrites the bf16 shadow copy the GEMMs read (work-list W7).

``state_dict``/``load_state_dict`` speak the two formats of SURVEY section 5.4: the stock
``torch.optim`` int-keyed layout with the reference's decay / no-decay groups (DDP) and the
FQN-keyed ``FSDP.optim_state_dict`` layout (FSDP), so checkpoints interchange.
"""
from __future__ import annotations

import math
from typing import Any, Dict, List, Optional

import torch

from ..parallel.flat import FlatParamStore


def adamw_reference_(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, decay_mask: torch.Tensor,
                     step: int, lr: float, beta1: float, beta2: float, eps: float, weight_decay: float,
                     grad_scale: float = 1.0, shadow: Optional[torch.Tensor] = None) -> None:
    """Plain-PyTorch statement of the kernel (oracle for tests, and the CPU path)."""
    g = g * grad_scale
    p.mul_(1.0 - lr * weight_decay * decay_mask)
    m.mul_(beta1).add_(g, alpha=1.0 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1.0 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)
    if shadow is not None:
        shadow.copy_(p)


class FlatAdamW:
    """Optimizer facade with the ``torch.optim`` surface the trainers touch
    (``param_groups[i]["lr"]``, ``step``, ``zero_grad``, ``state_dict``, ``load_state_dict``)."""

    def __init__(self, store: FlatParamStore, lr: float, betas=(0.9, 0.95), eps: float = 1e-8,
                 weight_decay: float = 0.1, ranges: Optional[List] = None, decay_norms: bool = False,
                 layout: str = "ddp"):
        self.store = store
        # flat element ranges this rank owns (sharded engines pass their slices); state is stored
        # back to back in range order
        self.ranges = [(0, store.total)] if ranges is None else [(int(a), int(b)) for a, b in ranges]
        self.state_offsets = []
        n = 0
        for a, b in self.ranges:
            self.state_offsets.append(n)
            n += b - a
        self.layout = layout            # "ddp" (two groups, int keys) or "fsdp" (one group, FQN keys)
        self.decay_norms = decay_norms  # reference FSDP decays every tensor (fsdp_trainer.py:338-343)
        dev = store.device
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self.step_count = 0
        # device scalar multiplied into every gradient (clip coefficient x 1/world, written by the comm kernels)
        self.grad_scale = torch.ones(1, dtype=torch.float32, device=dev)
        d_idx, nd_idx = store.group_indices()
        common = dict(lr=lr, betas=tuple(betas), eps=eps, amsgrad=False, maximize=False, foreach=None,
                      capturable=False, differentiable=False, fused=bool(dev.type == "cuda"),
                      decoupled_weight_decay=True)
        if layout == "ddp":
            self.param_groups: List[Dict[str, Any]] = [
                dict(common, weight_decay=weight_decay, params=list(range(len(d_idx)))),
                dict(common, weight_decay=0.0, params=list(range(len(d_idx), len(d_idx) + len(nd_idx)))),
            ]
            self._order = d_idx + nd_idx           # optimizer index -> slot index
        else:
            self.param_groups = [dict(common, weight_decay=weight_decay, fused=None,
                                      params=[s.name for s in store.slots])]
            self._order = [s.index for s in store.slots]
        if decay_norms:
            self._decay = torch.ones_like(store.decay_chunks)
        else:
            self._decay = store.decay_chunks

    # ---- hot path -----------------------------------------------------------------------------
    def step(self, grad_scale: Optional[torch.Tensor] = None) -> None:
        from . import kernels

        g0 = self.param_groups[0]
        lr, (b1, b2), eps, wd = g0["lr"], g0["betas"], g0["eps"], g0["weight_decay"]
        self.step_count += 1
        st = self.store
        scale = self.grad_scale if grad_scale is None else grad_scale
        for (lo, hi), off in zip(self.ranges, self.state_offsets):
            if hi <= lo:
                continue
            m, v = self.exp_avg[off: off + hi - lo], self.exp_avg_sq[off: off + hi - lo]
            if st.device.type == "cuda":
                kernels.adamw_step(st.master, st.grad, m, v, st.shadow, self._decay,
                                   lo, hi, self.step_count, lr, b1, b2, eps, wd, scale)
            else:
                from ..parallel.flat import ALIGN

                mask = self._decay.to(torch.float32).repeat_interleave(ALIGN)[lo:hi]
                adamw_reference_(st.master[lo:hi], st.grad[lo:hi], m, v, mask,
                                 self.step_count, lr, b1, b2, eps, wd, float(scale.item()),
                                 None if st.shadow is None else st.shadow[lo:hi])

    def zero_grad(self, set_to_none: bool = False) -> None:
        self.store.zero_grad()

    # ---- state dict --------------------------------------------------------------------------------
    # Checkpoints always carry the FULL (unsharded) state in the reference's format.  A sharded
    # optimizer is serialised by first assembling full flat exp_avg / exp_avg_sq buffers
    # (``NVLinkGradSync`` / ``ShardedGPT`` gather them) and calling ``state_dict(full_m, full_v)``.
    @property
    def is_full(self) -> bool:
        return self.ranges == [(0, self.store.total)]

    def state_dict(self, full_exp_avg: Optional[torch.Tensor] = None,
                   full_exp_avg_sq: Optional[torch.Tensor] = None) -> Dict[str, Any]:
        if full_exp_avg is None:
            assert self.is_full, "gather the sharded state first"
            full_exp_avg, full_exp_avg_sq = self.exp_avg, self.exp_avg_sq
        slots = self.store.slots
        state: Dict[Any, Any] = {}
        if self.step_count > 0:
            for k, si in enumerate(self._order):
                sl = slots[si]
                a, b = sl.offset, sl.offset + sl.numel
                state[k if self.layout == "ddp" else sl.name] = {
                    "step": torch.tensor(float(self.step_count)),
                    "exp_avg": full_exp_avg[a:b].view(sl.shape).clone(),
                    "exp_avg_sq": full_exp_avg_sq[a:b].view(sl.shape).clone(),
                }
        return {"state": state, "param_groups": [dict(g) for g in self.param_groups]}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        """Accepts either key style (int-keyed DDP layout or FQN-keyed FSDP layout); a sharded
        optimizer keeps only the elements inside its owned ranges."""
        slots = self.store.slots
        state = sd.get("state", {})
        name_to_slot = {s.name: s for s in slots}
        steps = set()
        for key, st in state.items():
            if isinstance(key, str):
                slot = name_to_slot[key]
            else:
                slot = slots[self._order[int(key)]] if self.layout == "ddp" else slots[int(key)]
            src_m, src_v = st["exp_avg"].reshape(-1), st["exp_avg_sq"].reshape(-1)
            for (lo, hi), off in zip(self.ranges, self.state_offsets):
                a, b = max(lo, slot.offset), min(hi, slot.offset + slot.numel)
                if a < b:
                    self.exp_avg[off + a - lo: off + b - lo].copy_(src_m[a - slot.offset: b - slot.offset])
                    self.exp_avg_sq[off + a - lo: off + b - lo].copy_(src_v[a - slot.offset: b - slot.offset])
            steps.add(int(float(st["step"])))
        self.step_count = max(steps) if steps else 0
        for g, src in zip(self.param_groups, sd.get("param_groups", [])):
            for k in ("lr", "betas", "eps", "weight_decay"):
                if k in src:
                    g[k] = tuple(src[k]) if k == "betas" else src[k]
