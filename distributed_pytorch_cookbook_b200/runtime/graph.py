"""Whole-step CUDA-graph capture (fwd + bwd + optimizer in one launch).

This replaces the reference's ``torch.compile`` (main-single.py:38-39) as the way to get
rid of per-kernel launch cost: at the default config a step is ~2 ms of GPU work spread
over a few hundred small kernels, so the host would otherwise be the bottleneck.  The first
``warmup`` calls run eagerly (they are real training steps); the next call is captured and
from then on every call is ``copy inputs -> replay``.
"""
from __future__ import annotations

from typing import Callable, List, Optional

import torch


class GraphedStep:
    def __init__(self, fn: Callable[..., torch.Tensor], warmup: int = 3):
        self.fn, self.warmup = fn, warmup
        self.calls = 0
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.static_in: List[torch.Tensor] = []
        self.static_out = None
        self.side = None
        self.native_calls_per_replay = 0

    def __call__(self, *tensors: torch.Tensor):
        if self.graph is None:
            if self.side is None:
                self.side = torch.cuda.Stream()
            if self.calls < self.warmup:
                self.calls += 1
                self.side.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(self.side):
                    out = self.fn(*tensors)
                torch.cuda.current_stream().wait_stream(self.side)
                return out
            self.static_in = [t.clone() for t in tensors]
            torch.cuda.synchronize()
            from ..ops import _lib
            n0 = _lib.launch_count()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph, stream=self.side):
                self.static_out = self.fn(*self.static_in)
            self.native_calls_per_replay = _lib.launch_count() - n0
        for s, t in zip(self.static_in, tensors):
            s.copy_(t, non_blocking=True)
        self.graph.replay()
        self.calls += 1
        out = self.static_out
        if isinstance(out, torch.Tensor):
            return out.clone()
        return tuple(o.clone() if isinstance(o, torch.Tensor) else o for o in out)
