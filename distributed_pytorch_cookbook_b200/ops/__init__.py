"""Op layer: one functional entry point per hot op, two back-ends each.

* ``*_reference`` - plain PyTorch math (runs anywhere, fp32 or whatever dtype it is
  given); it is the CPU path, the ``--disable_amp`` path and the test oracle.
* native - ``torch.autograd.Function`` wrappers (``ops.native``) around the sm_100a
  kernels of ``csrc/`` reached through the C ABI of ``include/b200gpt.h``.

``use_native(t)`` is the single dispatch predicate: CUDA tensor + library loaded +
bf16 autocast active (the reference expresses mixed precision with
``torch.autocast("cuda", bfloat16)``, main-single.py:88-90; we honour the same switch).
There is no silent fallback on a GPU box: if a CUDA tensor shows up under bf16
autocast and the library cannot be loaded, ``use_native`` raises.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.nn.functional as F

from . import _lib

_FORCE_REFERENCE = False


class force_reference:
    """Context manager: run the PyTorch reference math even on CUDA (used by tests)."""

    def __enter__(self):
        global _FORCE_REFERENCE
        self._old = _FORCE_REFERENCE
        _FORCE_REFERENCE = True

    def __exit__(self, *exc):
        global _FORCE_REFERENCE
        _FORCE_REFERENCE = self._old


def native_available() -> bool:
    return _lib.available()


def use_native(t: torch.Tensor) -> bool:
    if _FORCE_REFERENCE or not t.is_cuda:
        return False
    if not (torch.is_autocast_enabled("cuda") and torch.get_autocast_dtype("cuda") == torch.bfloat16):
        return False
    if not _lib.available():
        if os.environ.get("B200GPT_ALLOW_FALLBACK") == "1":
            return False
        raise RuntimeError(
            "CUDA tensor under bf16 autocast but the sm_100a library is not loaded: "
            f"{_lib.load_error()} (build it with `python -m csrc.build`; set "
            "B200GPT_ALLOW_FALLBACK=1 to run the PyTorch reference math instead)")
    return True


# ----------------------------------------------------------------------------------------
# reference math
# ----------------------------------------------------------------------------------------
def attention_reference(q, k, v, scale: float, key_mask: Optional[torch.Tensor] = None):
    """softmax(q k^T * scale + causal(-1e9) [key-pad -> finfo.min]) v, reference gpt.py:79-99.

    q, k, v: [N, h, S, d].  key_mask: bool [N, S], True = padding.
    """
    S = q.shape[-2]
    logits = (q @ k.transpose(-1, -2)) * scale
    causal = torch.ones(S, S, device=q.device, dtype=torch.bool).tril_()
    logits = logits + torch.where(causal, 0.0, -1e9).to(logits.dtype)
    if key_mask is not None:
        logits = logits.masked_fill(key_mask[:, None, None, :], torch.finfo(logits.dtype).min)
    return F.softmax(logits, dim=-1) @ v


def lmhead_cross_entropy_reference(x, weight, targets, ignore_index: int = -100):
    """logits = x W^T; mean CE over targets != ignore_index; argmax accuracy counts.

    Reference: lm_head gpt.py:231 + F.cross_entropy main-single.py:95-96 + accuracy :127-131.
    """
    logits = F.linear(x, weight if weight.dtype == x.dtype else weight.to(x.dtype))
    logits = logits.reshape(-1, logits.shape[-1])
    t = targets.reshape(-1)
    loss = F.cross_entropy(logits.float(), t, ignore_index=ignore_index)
    valid = t != ignore_index
    n_valid = valid.sum()
    n_correct = ((logits.argmax(dim=-1) == t) & valid).sum()
    return loss, n_valid, n_correct


# ----------------------------------------------------------------------------------------
# dispatching entry points
# ----------------------------------------------------------------------------------------
def lmhead_cross_entropy(x, weight, targets, ignore_index: int = -100, grad_scale: float = 1.0):
    """``grad_scale`` is the factor the caller will apply to the loss before ``backward()`` (1/M
    for M pipeline micro-batches); the native op bakes it into the gradients it produces in the
    forward pass, the PyTorch path gets it through autograd."""
    if use_native(x):
        from . import native
        return native.lmhead_cross_entropy(x, weight, targets, ignore_index, grad_scale)
    return lmhead_cross_entropy_reference(x, weight, targets, ignore_index)


def linear(x, weight, bias=None, activation: Optional[str] = None, residual=None,
           out_fp32: bool = False, handoff: str = "stream"):
    from . import native
    return native.linear(x, weight, bias, activation, residual, out_fp32, handoff)


def mlp(x, w1, b1, w2, b2, residual=None, activation: str = "relu"):
    from . import native
    return native.mlp(x, w1, b1, w2, b2, residual, activation)


def flush_wgrads() -> None:
    """Launch the weight-gradient GEMMs the native backward has queued (see ``native.wgrad``).  Every backward pass
    does this by itself when it ends; only code that reads ``.grad`` in the middle of a backward pass needs it."""
    from . import native
    native.flush_wgrads()


def linear_qkv(x, wq, wk, wv):
    from . import native
    return native.linear_qkv(x, wq, wk, wv)


def attention(qkv, heads: int, head_dim: int, scale: float, key_mask=None):
    from . import native
    return native.attention(qkv, heads, head_dim, scale, key_mask)


def layernorm(x, weight, bias, eps: float):
    from . import native
    return native.layernorm(x, weight, bias, eps)


def layernorm_residual(x, weight, bias, eps: float):
    from . import native
    return native.layernorm_residual(x, weight, bias, eps)


def embedding(input_ids, position_ids, tok_w, pos_w):
    from . import native
    return native.embedding(input_ids, position_ids, tok_w, pos_w)
