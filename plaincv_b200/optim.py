"""The step after the hot path: `apply_grads` of train_lm.py:316-353 for the reference's default
optimizer — `_clip_grads` (train_lm.py:173-178) followed by `optax.adamw(lr, b1, b2, eps, weight_decay)` without a decay
mask (optim/factory.py:193-205) — as two kernels over the model's flat buffers: a sum of squares of the gradients
(block partials, reproducible) and one pass that clips, updates m / v / the fp32 masters and re-writes the bf16 compute
copies (so `sync_compute_weights` is not needed after it).  The other optimizers of optim/factory.py are out of scope.

Every rank holds the same averaged gradients (Engine.compute_grads / train_step), so every rank applies the same
update to its replica, as the reference's pmap'd apply_grads does.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops


class AdamW:
    """optax.adamw + the reference's global-norm clip over a plaincv_b200.Transformer.

    `lr` may be overridden per step (`step(lr=...)`): the reference's warm-up / cooldown schedule
    (utils/schedules in the reference, out of scope) is a function of the step count evaluated by the caller."""

    def __init__(self, model, lr: float, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8,
                 weight_decay: float = 0.0, grad_clip: Optional[float] = None):
        if not model.flat_master.is_cuda:
            raise RuntimeError("plaincv_b200.optim.AdamW needs the model on a CUDA device (there is no CPU fallback)")
        if grad_clip is not None and grad_clip <= 0:
            raise ValueError("grad_clip must be positive (or None)")
        self.model = model
        self.lr, self.b1, self.b2, self.eps, self.wd = float(lr), float(beta1), float(beta2), float(eps), float(weight_decay)
        self.grad_clip = None if grad_clip is None else float(grad_clip)
        dev = model.flat_master.device
        self.m = torch.zeros_like(model.flat_master)
        self.v = torch.zeros_like(model.flat_master)
        self.count = 0  # optax's `count`: number of updates applied
        self._partials = torch.zeros(ops._NUM_SMS * 4, dtype=torch.float32, device=dev)
        self.grad_norm = torch.zeros(1, dtype=torch.float32, device=dev)  # global norm seen by the last step

    @classmethod
    def from_cfg(cls, model, cfg) -> "AdamW":
        """The reference's YAML keys (config/tr_420M_x8gpu.yaml:43-47; optim/factory.py:193-205)."""
        get = cfg.get if isinstance(cfg, dict) else (lambda k, d=None: getattr(cfg, k, d))
        name = str(get("optim", "adamw")).lower()
        if name not in ("adam", "adamw"):
            raise NotImplementedError(f"optimizer {name!r}: only adam / adamw of optim/factory.py are built here")
        return cls(model, lr=float(get("lr")), beta1=get("beta1", 0.9), beta2=get("beta2", 0.999), eps=get("eps", 1e-8),
                   weight_decay=get("weight_decay", 0.0), grad_clip=get("grad_clip", None))

    @torch.no_grad()
    def step(self, lr: Optional[float] = None) -> None:
        """One update from model.flat_grad (already averaged over micro-batches and ranks)."""
        m = self.model
        self.count += 1
        ops.sumsq_partials(m.flat_grad, self._partials)
        ops.adamw_step(m.flat_master, m.flat_grad, self.m, self.v, m.flat_bf16, self._partials, self.grad_norm,
                       self.grad_clip if self.grad_clip is not None else 0.0, self.lr if lr is None else lr,
                       self.b1, self.b2, self.eps, self.wd, self.count)

    def state_dict(self) -> dict:
        return {"m": self.m, "v": self.v, "count": self.count}

    def load_state_dict(self, sd: dict) -> None:
        self.m.copy_(sd["m"])
        self.v.copy_(sd["v"])
        self.count = int(sd["count"])
