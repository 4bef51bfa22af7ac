"""Flat parameter / gradient arenas.

B200-first memory layout: all fp32 master weights live in ONE flat buffer, all fp32
gradients in ONE flat buffer of the same geometry (so a DDP bucket is just a slice, and the
slice can live in NVLink-mapped symmetric memory), the AdamW moments in two more, and the
bf16 compute copies the tensor-core kernels read in a fifth.  ``nn.Parameter`` objects keep
their identity and names (the reference's state_dict contract) - their
``.data``/``.grad`` are re-pointed at views of the arenas.  One fused AdamW launch then
updates the whole model; one memset clears every gradient.
"""
from __future__ import annotations

from typing import Callable, List, Optional

import torch

ALIGN = 128  # elements; keeps every bf16 view 256-byte aligned (TMA needs 16 B)


def _round_up(n: int, m: int) -> int:
    return (n + m - 1) // m * m


class ParamArena:
    def __init__(self, params: List[torch.nn.Parameter], device, with_bf16: bool = True,
                 with_moments: bool = True,
                 grad_alloc: Optional[Callable[[int], torch.Tensor]] = None,
                 pad_total_to: int = ALIGN):
        self.params = list(params)
        self.device = torch.device(device)
        self.offsets, off = [], 0
        for p in self.params:
            self.offsets.append(off)
            off += _round_up(p.numel(), ALIGN)
        self.total = _round_up(off, pad_total_to)
        dev = self.device
        self.fp32 = torch.zeros(self.total, dtype=torch.float32, device=dev)
        self.grad = grad_alloc(self.total) if grad_alloc is not None else \
            torch.zeros(self.total, dtype=torch.float32, device=dev)
        self.bf16 = torch.zeros(self.total, dtype=torch.bfloat16, device=dev) if with_bf16 else None
        self.exp_avg = torch.zeros(self.total, dtype=torch.float32, device=dev) if with_moments else None
        self.exp_avg_sq = torch.zeros(self.total, dtype=torch.float32, device=dev) if with_moments else None
        with torch.no_grad():
            for p, o in zip(self.params, self.offsets):
                n = p.numel()
                view = self.fp32[o:o + n].view(p.shape)
                view.copy_(p.data.to(dev))
                p.data = view
                p.grad = self.grad[o:o + n].view(p.shape)
                p._b200_arena = self
                p._b200_offset = o
        self.refresh_bf16()

    def refresh_bf16(self):
        """(Re)derive the bf16 compute copies from the fp32 masters."""
        if self.bf16 is None:
            return
        with torch.no_grad():
            self.bf16.copy_(self.fp32)
            for p, o in zip(self.params, self.offsets):
                p._b200_bf16 = self.bf16[o:o + p.numel()].view(p.shape)
                p._b200_bf16_version = p._version

    def zero_grad(self):
        self.grad.zero_()

    def slice_of(self, p) -> slice:
        o = p._b200_offset
        return slice(o, o + p.numel())
