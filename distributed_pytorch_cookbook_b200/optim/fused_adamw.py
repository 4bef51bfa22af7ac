"""AdamW with the reference's hyper-parameters, fused over a flat arena.

Reference: ``torch.optim.AdamW(model.parameters(), lr)`` (main-single.py:42) => betas
(0.9, 0.999), eps 1e-8, weight_decay 0.01 on *every* parameter, no amsgrad; wrapped in a
GradScaler (main-single.py:78,99-101) that for bf16 only adds an unscale pass, an inf check
and a host sync per step.  Here ONE kernel launch (``b200_adamw``, csrc/elementwise.cu)
updates all parameters, writes the bf16 compute copies in the same pass, reads lr/step from
device memory (so it replays inside a CUDA graph).  Non-finite gradients: the reference's GradScaler
skips the *whole* optimizer step when any gradient element is inf/nan, at the price of a host sync per
step; this kernel skips *per element* - that parameter and its moments keep their values, every finite
element still updates, the step counter still advances - and counts the skipped elements in
``hyper[2]`` (``skipped_elements()``), with no host round trip.  A deliberate deviation: under bf16
autocast (no loss scaling needed) non-finite gradients mean the run is already in trouble, and the
counter makes that visible without stalling every healthy step.
On CPU (or when the library is absent) the same maths runs through ``torch.optim.AdamW``.
"""
from __future__ import annotations

from typing import Optional

import torch

from ..ops import _lib
from ..runtime.arena import ParamArena


class FusedAdamW:
    def __init__(self, params, lr: float = 1e-4, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, device=None, arena: Optional[ParamArena] = None,
                 grad_alloc=None):
        self.params = [p for p in params if p.requires_grad]
        self.lr, self.betas, self.eps, self.weight_decay = lr, betas, eps, weight_decay
        device = torch.device(device if device is not None else self.params[0].device)
        self.native = device.type == "cuda" and _lib.available()
        if self.native:
            self.arena = arena if arena is not None else ParamArena(self.params, device, grad_alloc=grad_alloc)
            # hyper[0] = lr, hyper[1] = step (float, exact up to 2^24 steps), hyper[2] = gradient elements skipped so far
            self.hyper = torch.tensor([lr, 0.0, 0.0, 0.0], dtype=torch.float32, device=device)
            self.torch_opt = None
        else:
            self.arena = arena          # optional: flat fp32 params/grads (DDP buckets) on any device
            self.torch_opt = torch.optim.AdamW(self.params, lr=lr, betas=betas, eps=eps,
                                               weight_decay=weight_decay)

    # -- torch.optim-like surface ---------------------------------------------------------
    def zero_grad(self, set_to_none: bool = False):
        if self.arena is not None:
            self.arena.zero_grad()
        else:
            self.torch_opt.zero_grad(set_to_none=set_to_none)

    def set_lr(self, lr: float):
        self.lr = lr
        if self.native:
            self.hyper[0] = lr
        else:
            for g in self.torch_opt.param_groups:
                g["lr"] = lr

    @torch.no_grad()
    def step(self, lo: int = 0, hi: Optional[int] = None, tick: bool = True):
        """Update elements [lo, hi) of the arena (whole arena by default)."""
        if not self.native:
            self.torch_opt.step()
            return
        a = self.arena
        hi = a.total if hi is None else hi
        s = _lib.stream()
        if tick:
            _lib.call("b200_adamw_tick", self.hyper.data_ptr(), s)
        _lib.call("b200_adamw", a.fp32.data_ptr() + 4 * lo, a.grad.data_ptr() + 4 * lo,
                  a.exp_avg.data_ptr() + 4 * lo, a.exp_avg_sq.data_ptr() + 4 * lo,
                  a.bf16.data_ptr() + 2 * lo, hi - lo, self.hyper.data_ptr(),
                  self.betas[0], self.betas[1], self.eps, self.weight_decay, s)

    def skipped_elements(self) -> int:
        """Gradient elements the kernel has skipped as non-finite since the start (host sync)."""
        return int(self.hyper[2]) if self.native else 0

    def state_dict(self):
        if self.native:
            a = self.arena
            return {"native": True, "hyper": self.hyper.cpu(), "exp_avg": a.exp_avg.cpu(),
                    "exp_avg_sq": a.exp_avg_sq.cpu()}
        return {"native": False, "torch": self.torch_opt.state_dict()}

    def load_state_dict(self, sd):
        """Moments and step counter come from ``sd``; the learning rate stays the one this run was started with."""
        if self.native and sd.get("native"):
            self.hyper.copy_(sd["hyper"])
            self.hyper[0] = self.lr
            self.arena.exp_avg.copy_(sd["exp_avg"])
            self.arena.exp_avg_sq.copy_(sd["exp_avg_sq"])
        elif not self.native and not sd.get("native"):
            self.torch_opt.load_state_dict(sd["torch"])
            for g in self.torch_opt.param_groups:
                g["lr"] = self.lr
        else:
            raise ValueError("optimizer state was saved by the other back-end")


def build_optimizer(model, lr: float, device=None, **kw) -> FusedAdamW:
    return FusedAdamW(model.parameters(), lr=lr, device=device, **kw)
