"""NVTX ranges around the phases of a training step (``--profile``).

``enable()`` is called by the Trainer when ``--profile`` is given; ``phase(name)`` is a context manager the
strategies put around forward / backward / gradient sync / optimizer.  Disabled (the default) it costs one
attribute read.  With ``--profile`` the strategies also run the step eagerly instead of replaying a CUDA
graph, so the ranges bracket the kernels they name in an ``nsys`` / ``ncu --nvtx`` timeline.
"""
from __future__ import annotations

from contextlib import contextmanager

import torch

_ON = False


def enable(on: bool = True):
    global _ON
    _ON = bool(on) and torch.cuda.is_available()


def enabled() -> bool:
    return _ON


@contextmanager
def phase(name: str):
    if not _ON:
        yield
        return
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()
