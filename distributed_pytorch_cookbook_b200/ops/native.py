"""autograd wrappers around the sm_100a kernels (C ABI in include/b200gpt.h).

Conventions shared by all wrappers:

* activations between kernels are bf16, the residual stream / LayerNorm input is fp32
  (the reference's autocast contract);
* weights are fp32 ``nn.Parameter`` masters; kernels read a cached bf16 copy
  (``bf16_of``) that the fused AdamW keeps up to date, so no per-step cast exists;
* weight/bias gradients are accumulated by the kernels *directly into* ``param.grad``
  (a view of the flat gradient arena when one exists) and the autograd edge gets ``None``:
  no AccumulateGrad copy, and a DDP bucket is ready the moment the kernel is queued.
  ``grad_ready_hooks`` are called with the parameter right after that launch.
"""
from __future__ import annotations

from typing import Callable, List, Optional

import torch

from . import _lib
from ._lib import call, ptr, stream

BF16 = torch.bfloat16
grad_ready_hooks: List[Callable[[torch.nn.Parameter], None]] = []
# per-parameter override of where the bf16 weights come from (FSDP installs gathered buffers)
import ctypes as _ctypes
import os as _os
# rows of bf16 logits in flight (one reused buffer of rows x vocab; 16384 rows = 1.6 GB at vocab 50257).  Bigger
# chunks mean fewer, larger GEMMs and fewer read-modify-write passes over dW: 4.93 -> 4.86 ms/step from 4096 to 16384.
_LM_CHUNK_ROWS = int(_os.environ.get("B200GPT_LM_CHUNK", "16384"))
_LM_SUB_ROWS = int(_os.environ.get("B200GPT_LM_SUB", "0"))     # rows per logits-GEMM + CE sub-chunk (0 = whole chunk)


_SCRATCH = {}
SCRATCH_COLS = 12288   # 3 x the widest LayerNorm row (dgamma, dbeta and the handed-off bias gradient)


def colsum_scratch(device):
    """Per-device scratch of the column-reduction kernels (bias / gamma / beta gradients):
    16 replica rows x 8192 floats + a ticket counter, zero between launches (the kernels clean up
    after themselves).  Shared by every launch on the device's compute stream - which is the only
    stream these kernels run on - so launches never overlap."""
    key = (device.type, device.index)
    s = _SCRATCH.get(key)
    if s is None:
        s = (torch.zeros(16 * SCRATCH_COLS, dtype=torch.float32, device=device),
             torch.zeros(1, dtype=torch.int32, device=device))
        _SCRATCH[key] = s
    return s


def _notify(*params):
    if grad_ready_hooks:
        for p in params:
            if p is not None:
                for h in grad_ready_hooks:
                    h(p)


def bf16_of(p: torch.Tensor) -> torch.Tensor:
    """bf16 compute copy of an fp32 master weight (cached on the parameter)."""
    if p.dtype == BF16:
        return p
    cached = getattr(p, "_b200_bf16", None)
    if cached is not None and getattr(p, "_b200_bf16_version", -1) == p._version:
        return cached
    with torch.no_grad():
        if cached is None or cached.shape != p.shape or cached.device != p.device:
            cached = torch.empty(p.shape, dtype=BF16, device=p.device)
            p._b200_bf16 = cached
        call("b200_cast_bf16", p.data_ptr(), cached.data_ptr(), p.numel(), stream())
        p._b200_bf16_version = p._version
    return cached


def grad_of(p: torch.nn.Parameter) -> torch.Tensor:
    """fp32 gradient buffer the kernels accumulate into."""
    if p.grad is None:
        p.grad = torch.zeros_like(p, dtype=torch.float32)
    return p.grad


def _as_bf16_2d(x: torch.Tensor) -> torch.Tensor:
    x2 = x.reshape(-1, x.shape[-1])
    if x2.dtype != BF16:
        x2 = x2.to(BF16)
    return x2.contiguous()


def _tma_rows(t: torch.Tensor) -> torch.Tensor:
    """A 2-D tensor the TMA engine can address: unit column stride, 16-byte aligned base and row pitch.
    Anything else (e.g. bf16 rows of 50257 elements) is copied into rows padded to the next 16 bytes."""
    esz = t.element_size()
    if t.stride(1) == 1 and (t.stride(0) * esz) % 16 == 0 and t.data_ptr() % 16 == 0:
        return t
    per = 16 // esz
    ld = (t.shape[1] + per - 1) // per * per
    buf = torch.empty((t.shape[0], ld), dtype=t.dtype, device=t.device)
    out = buf[:, :t.shape[1]]
    out.copy_(t)
    return out


def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int, *,
         a_mn_major: bool = False, b_mn_major: bool = False, lda=None, ldb=None,
         bias=None, residual=None, act: int = 0, accumulate: bool = False,
         aux_out=None, mask_src=None, cfg: int = 0, ag=None):
    """out[M,N] (+)= op(A)[M,K] . op(B)[N,K]^T with a fused epilogue (csrc/gemm_tcgen05.cu).

    K-major operand: stored [rows, K] (row-major, ``ld`` = row stride).  MN-major operand:
    stored [K, rows] (i.e. the transpose is what sits in memory) - used by dgrad/wgrad so
    no tensor is ever transposed in memory.
    """
    lda = lda if lda is not None else a.stride(0)
    ldb = ldb if ldb is not None else b.stride(0)
    if ag is not None:   # FSDP: the all-gather of this weight is fused into the GEMM (csrc/gemm_tcgen05.cu)
        peers_dev, full_base, off, nbytes, shard_bytes, sync, sync_tiles = ag
        assert not a_mn_major and not b_mn_major and not accumulate and mask_src is None
        call("b200_ag_gemm_bf16", a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, lda, ldb, out.stride(0),
             ptr(bias), ptr(residual), residual.stride(0) if residual is not None else 0, act,
             int(out.dtype == torch.float32), ptr(aux_out), peers_dev, full_base, off, nbytes, shard_bytes, sync,
             sync_tiles, stream())
        return out
    call("b200_gemm_bf16", a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, lda, ldb,
         out.stride(0), int(a_mn_major), int(b_mn_major), ptr(bias), ptr(residual),
         residual.stride(0) if residual is not None else 0, act, int(out.dtype == torch.float32),
         int(accumulate), ptr(aux_out), ptr(mask_src), cfg, stream())
    return out


def fused_gather_args(weights):
    """FSDP hook: if these (adjacent) weights are sharded and their gathered bf16 copy is stale,
    return the arguments that make the consuming GEMM pull them from the peers itself, and mark
    them gathered.  ``None`` when there is nothing to do (not sharded / already resident)."""
    info = getattr(weights[0], "_b200_ag", None)
    if info is None:
        return None
    eng, grp, off0, _ = info
    if not eng.fused_ag or not grp.fused or all(getattr(w, "_b200_ag_version", -1) == eng.version for w in weights):
        return None
    last = weights[-1]._b200_ag
    end = last[2] + last[3]
    for w in weights:
        w._b200_ag_version = eng.version
    return (grp.peers_dev.data_ptr(), eng.full_bf16.data_ptr() + 2 * grp.base, off0, end - off0,
            2 * grp.shard_n, grp.ag_sync.data_ptr(), grp.ag_sync.numel() - 2)


# ----------------------------------------------------------------------------------------
# ----------------------------------------------------------------------------------------
# Grouped weight gradients.  A small dW GEMM (a few output tiles, K = tokens) costs ~13 us however little it computes:
# launch, tensor-memory allocation, pipeline fill and the split-K reduce-add tail dominate.  The dW GEMMs of one
# transformer block are therefore queued and launched together (``b200_gemm_wgrad_group``: one persistent grid, one
# K split).  The queue is flushed when it holds 4 problems, when the LayerNorm at a block's entry runs its backward
# (so under FSDP a unit's gradients are complete inside that unit's backward), and by an autograd-engine callback at
# the end of every backward pass; the gradient-ready notifications of the queued weights are issued at the flush.
_WGRAD_GROUP = (_os.environ.get("B200GPT_WGRAD_GROUP", "1") != "0" and _os.environ.get("B200GPT_GEMM_2SM", "1") != "0")
_WGRAD_MAX_TILES = 16          # defer only GEMMs that leave most of the GPU idle on their own
_wgrad_queue: list = []        # (dz, x, gw, n_out, k_in, tokens, lda, params)
_wgrad_cb_armed = [False]


def flush_wgrads() -> None:
    q = list(_wgrad_queue)
    _wgrad_queue.clear()
    if not q:
        return
    if len(q) == 1:
        dz, x, gw, n_out, k_in, tok, lda, params = q[0]
        gemm(dz, x, gw, n_out, k_in, tok, a_mn_major=True, b_mn_major=True, accumulate=True, lda=lda)
    else:
        n = len(q)
        vp, ci = _ctypes.c_void_p, _ctypes.c_int
        arr = lambda t, vals: _ctypes.cast((t * n)(*vals), vp)   # noqa: E731
        call("b200_gemm_wgrad_group", n, arr(vp, [e[0].data_ptr() for e in q]), arr(vp, [e[1].data_ptr() for e in q]),
             arr(vp, [e[2].data_ptr() for e in q]), arr(ci, [e[3] for e in q]), arr(ci, [e[4] for e in q]), q[0][5],
             arr(ci, [e[6] for e in q]), arr(ci, [e[1].stride(0) for e in q]), arr(ci, [e[2].stride(0) for e in q]),
             stream())
    for e in q:
        _notify(*e[7])


def _flush_wgrads_at_end() -> None:
    _wgrad_cb_armed[0] = False
    flush_wgrads()


def wgrad(dz: torch.Tensor, x: torch.Tensor, gw: torch.Tensor, n_out: int, k_in: int, tokens: int, params, lda=None):
    """gw[n_out, k_in] (fp32) += dz^T x, then ``_notify(*params)`` - deferred and grouped when the GEMM is small."""
    lda = lda if lda is not None else dz.stride(0)
    tiles = ((n_out + 255) // 256) * ((k_in + 255) // 256)
    if not _WGRAD_GROUP or tiles > _WGRAD_MAX_TILES or n_out <= 128 or gw.stride(0) % 4 != 0 or gw.data_ptr() % 16 != 0:
        gemm(dz, x, gw, n_out, k_in, tokens, a_mn_major=True, b_mn_major=True, accumulate=True, lda=lda)
        _notify(*params)
        return
    if _wgrad_queue and _wgrad_queue[0][5] != tokens:
        flush_wgrads()
    _wgrad_queue.append((dz, x, gw, n_out, k_in, tokens, lda, tuple(params)))
    if not _wgrad_cb_armed[0]:
        try:   # runs when the current backward pass is over, whoever started it
            torch.autograd.Variable._execution_engine.queue_callback(_flush_wgrads_at_end)
            _wgrad_cb_armed[0] = True
        except RuntimeError:      # not inside a backward pass: nothing would flush later
            flush_wgrads()
            return
    if len(_wgrad_queue) >= 4:
        flush_wgrads()


# ----------------------------------------------------------------------------------------
# LayerNorm-backward hand-off.  In a pre-LN block the residual stream x is written by the epilogue of
# a Linear / MLP (reference models/gpt.py:119-123: ``x = x + attn(...)``, ``x = x + ff(...)``) and read by exactly one
# node, the next LayerNorm.  That LayerNorm's backward produces dx - which is the producer's incoming gradient - so
# it also writes what the producer's backward would compute first: dz = bf16(dx * act'(.)) and the bias gradient's
# column sums (``b200_layernorm_bwd_fused``).  The producer then skips its own pass over dx.  The producer's autograd
# node carries ``fuse_info``; the LayerNorm leaves ``fused = (dz, dx.data_ptr())`` on it.
_LN_FUSE = _os.environ.get("B200GPT_LN_FUSE", "1") != "0"
_FUSE_SCOPE = "all"


def set_fuse_scope(scope: str) -> None:
    """"all": any producer; "block": only a producer inside the same transformer block as the LayerNorm (the
    attention out-projection -> ln2).  FSDP needs "block": a unit's gradient buffer is only valid during that unit's
    backward, and the MLP -> next block's ln1 hand-off would write the bias gradient of the previous unit early."""
    global _FUSE_SCOPE
    assert scope in ("all", "block", "off")
    _FUSE_SCOPE = scope


def _producer_of(x: torch.Tensor):
    """-> (autograd node of the layer that wrote x if the hand-off applies, else None;
           True if that layer sits in the same transformer block as the LayerNorm reading x)"""
    node = x.grad_fn if torch.is_grad_enabled() else None
    info = getattr(node, "fuse_info", None)
    mid_block = info is not None and info[0] == "block"
    if not _LN_FUSE or _FUSE_SCOPE == "off" or info is None or x.dtype != torch.float32:
        return None, mid_block
    if _FUSE_SCOPE == "block" and not mid_block:
        return None, mid_block
    if x.shape[-1] % 4 != 0 or not x.is_contiguous():
        return None, mid_block
    return node, mid_block


def _take_handoff(ctx, dy2: torch.Tensor):
    h = ctx.__dict__.pop("fused", None)
    if h is None:
        return None
    dz, ptr_dx = h
    if ptr_dx != dy2.data_ptr() or dz.shape != dy2.shape:
        raise RuntimeError("fused LayerNorm backward: the residual stream has a second consumer, so the gradient this "
                           "layer received is not the one the LayerNorm handed over (set B200GPT_LN_FUSE=0)")
    return dz


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, residual, act, out_fp32, params, grad_on, kind="stream"):
        lead = x.shape[:-1]
        x2 = _as_bf16_2d(x)
        M, K = x2.shape
        wb = bf16_of(weight)
        N = wb.shape[0]
        # rows are padded to 16 bytes when N is not (vocab 50257: reference gpt.py:231 returns [.., vocab] logits);
        # the caller sees the [:, :N] view
        per = 4 if out_fp32 else 8
        out = torch.empty((M, (N + per - 1) // per * per), dtype=torch.float32 if out_fp32 else BF16,
                          device=x.device)[:, :N]
        res2 = residual.reshape(M, N) if residual is not None else None
        need_aux = act != 0 and residual is not None and grad_on
        aux = torch.empty((M, N), dtype=BF16, device=x.device) if need_aux else None
        gemm(x2, wb, out, M, N, K, bias=bias, residual=res2, act=act, aux_out=aux, ag=fused_gather_args([weight]))
        ctx.act, ctx.params, ctx.has_res = act, params, residual is not None
        ctx.in_dtype = x.dtype
        y_act = None
        if act != 0:
            y_act = aux if need_aux else out
        ctx.save_for_backward(x2, wb, y_act)
        ctx.fuse_info = (kind, y_act, act, bias) if (residual is not None and out_fp32 and grad_on) else None
        return out.view(*lead, N)

    @staticmethod
    def backward(ctx, dy):
        x2, wb, y_act = ctx.saved_tensors
        weight, bias = ctx.params
        M, K = x2.shape
        N = wb.shape[0]
        dy2 = dy.reshape(M, N)
        dres = dy if ctx.has_res else None
        need_dz = ctx.act != 0 or dy2.dtype != BF16 or bias is not None
        dz = _take_handoff(ctx, dy2)
        if dz is not None:
            pass                                  # the LayerNorm that consumed our output already wrote dz and db
        elif need_dz:
            if not dy2.is_contiguous():
                dy2 = dy2.contiguous()
            dz = torch.empty((M, N), dtype=BF16, device=dy.device)
            dbias = grad_of(bias) if (bias is not None and bias.requires_grad) else None
            sc, cnt = colsum_scratch(dy.device)
            call("b200_relu_bwd", dy2.data_ptr(), ptr(y_act), dz.data_ptr(), ptr(dbias), M, N,
                 int(dy2.dtype == torch.float32), sc.data_ptr(), cnt.data_ptr(), stream())
        else:
            dz = dy2
        dz = _tma_rows(dz)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, K), dtype=BF16, device=dy.device)
            gemm(dz, wb, dx, M, K, N, b_mn_major=True)          # dX = dZ . W
            dx = dx.view(*dy.shape[:-1], K)
            if ctx.in_dtype != BF16:
                dx = dx.to(ctx.in_dtype)
        if weight.requires_grad:
            wgrad(dz, x2, grad_of(weight), N, K, M, (weight, bias))                        # dW += dZ^T . X
        else:
            _notify(weight, bias)
        return dx, None, None, dres, None, None, None, None, None


def linear(x, weight, bias=None, activation: Optional[str] = None, residual=None, out_fp32=False,
           handoff: str = "stream"):
    """``handoff``: where the LayerNorm that reads the residual output lives - "block" (same transformer block, e.g.
    attention out-projection -> ln2) or "stream" (the next block / the final norm); see set_fuse_scope."""
    act = {None: 0, "relu": 1}[activation]
    return _Linear.apply(x, weight, bias, residual, act, out_fp32, (weight, bias), torch.is_grad_enabled(), handoff)


# ----------------------------------------------------------------------------------------
ACT_CODES = {"relu": 1, "gelu": 2, "gelu_tanh": 3}


class _MLP(torch.autograd.Function):
    """act(act(x W1^T + b1) W2^T + b2) [+ residual] as one autograd node (reference gpt.py:33-41; the
    activation is a constructor argument there, :14-15 - ReLU, GELU erf form and GELU tanh form are fused).

    Forward is two epilogue-fused GEMMs.  Backward fuses the first activation's derivative into the dgrad
    GEMM that produces its input gradient (``mask_src`` epilogue: dZ1 = (dZ2 W2) * act'(.)) together with
    the bias gradient db1 (column sums by a warp butterfly in the same epilogue), so the [tokens, 4*dim]
    activation gradient is written once, already masked, and never re-read.  ReLU's derivative is read off
    the activated output; GELU's needs the pre-activation, which the forward GEMM then stores next to its
    output (``aux_out``).
    """

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, residual, params, grad_on, act):
        lead = x.shape[:-1]
        x2 = _as_bf16_2d(x)
        M, K = x2.shape
        w1b, w2b = bf16_of(w1), bf16_of(w2)
        F1, N = w1b.shape[0], w2b.shape[0]
        gelu = act >= 2
        u = torch.empty((M, F1), dtype=BF16, device=x.device)
        z1 = torch.empty((M, F1), dtype=BF16, device=x.device) if (gelu and grad_on) else None
        gemm(x2, w1b, u, M, F1, K, bias=b1, act=act, aux_out=z1, ag=fused_gather_args([w1]))
        res2 = residual.reshape(M, N) if residual is not None else None
        out = torch.empty((M, N), dtype=torch.float32 if residual is not None else BF16, device=x.device)
        # second activation's backward operand: ReLU -> the activated output itself (bf16 copy when the output is the
        # fp32 residual sum), GELU -> the pre-activation
        need_aux = grad_on and (residual is not None or gelu)
        aux = torch.empty((M, N), dtype=BF16, device=x.device) if need_aux else None
        gemm(u, w2b, out, M, N, F1, bias=b2, residual=res2, act=act, aux_out=aux, ag=fused_gather_args([w2]))
        ctx.params, ctx.has_res, ctx.in_dtype, ctx.act = params, residual is not None, x.dtype, act
        ctx.save_for_backward(x2, w1b, w2b, u, aux if aux is not None else out, z1)
        ctx.fuse_info = ("stream", aux, act, b2) if (residual is not None and grad_on and aux is not None) else None
        return out.view(*lead, N)

    @staticmethod
    def backward(ctx, dy):
        x2, w1b, w2b, u, y_act, z1 = ctx.saved_tensors
        w1, b1, w2, b2 = ctx.params
        act = ctx.act
        M, K = x2.shape
        F1, N = w1b.shape[0], w2b.shape[0]
        dy2 = dy.reshape(M, N)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        sc, cnt = colsum_scratch(dy.device)
        dz2 = _take_handoff(ctx, dy2)
        if dz2 is None:
            dz2 = torch.empty((M, N), dtype=BF16, device=dy.device)
            call("b200_act_bwd", dy2.data_ptr(), y_act.data_ptr(), dz2.data_ptr(), grad_of(b2).data_ptr(), M, N,
                 int(dy2.dtype == torch.float32), act, sc.data_ptr(), cnt.data_ptr(), stream())
        wgrad(dz2, u, grad_of(w2), N, F1, M, (w2, b2))                                             # dW2 += dZ2^T u
        dz1 = torch.empty((M, F1), dtype=BF16, device=dy.device)
        # dZ1 = (dZ2 W2) * act'(.), with db1 += column sums of dZ1 folded into the same epilogue
        call("b200_gemm_dact_colsum_bf16", dz2.data_ptr(), w2b.data_ptr(), dz1.data_ptr(), M, F1, N, dz2.stride(0),
             w2b.stride(0), dz1.stride(0), 1, (z1 if act >= 2 else u).data_ptr(), grad_of(b1).data_ptr(), act, 0, stream())
        wgrad(dz1, x2, grad_of(w1), F1, K, M, (w1, b1))                                            # dW1 += dZ1^T x
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, K), dtype=BF16, device=dy.device)
            gemm(dz1, w1b, dx, M, K, F1, b_mn_major=True)                                         # dX = dZ1 W1
            dx = dx.view(*dy.shape[:-1], K)
            if ctx.in_dtype != BF16:
                dx = dx.to(ctx.in_dtype)
        return dx, None, None, None, None, (dy if ctx.has_res else None), None, None, None


def mlp(x, w1, b1, w2, b2, residual=None, activation: str = "relu"):
    return _MLP.apply(x, w1, b1, w2, b2, residual, (w1, b1, w2, b2), torch.is_grad_enabled(), ACT_CODES[activation])


# ----------------------------------------------------------------------------------------
def _packed(ts) -> Optional[torch.Tensor]:
    """If the tensors are adjacent rows of one buffer return the packed [sum_rows, K] view."""
    t0 = ts[0]
    rows = 0
    base = t0.untyped_storage().data_ptr()
    for t in ts:
        if (t.untyped_storage().data_ptr() != base or not t.is_contiguous()
                or t.data_ptr() != t0.data_ptr() + rows * t0.shape[1] * t0.element_size()):
            return None
        rows += t.shape[0]
    return torch.as_strided(t0, (rows, t0.shape[1]), (t0.shape[1], 1))


class _LinearQKV(torch.autograd.Function):
    """q,k,v projections as one GEMM over the three (adjacent) weight matrices.

    The weights stay three parameters (state_dict contract); in the arena they are adjacent,
    so the bf16 copies form one [3*inner, dim] operand and the three fp32 gradients one
    [3*inner, dim] output.  Otherwise they are concatenated on the fly.
    """

    @staticmethod
    def forward(ctx, x, wq, wk, wv, params):
        lead = x.shape[:-1]
        x2 = _as_bf16_2d(x)
        M, K = x2.shape
        ws = [bf16_of(w) for w in (wq, wk, wv)]
        wp = _packed(ws)
        ag = fused_gather_args([wq, wk, wv]) if wp is not None else None
        if wp is None:
            for w in (wq, wk, wv):      # not adjacent: gather each on its own before concatenating
                a1 = fused_gather_args([w])
                if a1 is not None:
                    raise RuntimeError("sharded q/k/v weights must be adjacent in their shard group")
            wp = torch.cat(ws, dim=0)
        N = wp.shape[0]
        out = torch.empty((M, N), dtype=BF16, device=x.device)
        gemm(x2, wp, out, M, N, K, ag=ag)
        ctx.params = params
        ctx.in_dtype = x.dtype
        ctx.save_for_backward(x2, wp)
        return out.view(*lead, N)

    @staticmethod
    def backward(ctx, dy):
        x2, wp = ctx.saved_tensors
        M, K = x2.shape
        N = wp.shape[0]
        dz = dy.reshape(M, N)
        if not dz.is_contiguous():
            dz = dz.contiguous()
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty((M, K), dtype=BF16, device=dy.device)
            gemm(dz, wp, dx, M, K, N, b_mn_major=True)
            dx = dx.view(*dy.shape[:-1], K)
            if ctx.in_dtype != BF16:
                dx = dx.to(ctx.in_dtype)
        params = ctx.params
        grads = [grad_of(p) for p in params]
        gp = _packed(grads)
        if gp is not None:
            wgrad(dz, x2, gp, N, K, M, params)
            return dx, None, None, None, None
        else:
            r = 0
            for prm, g in zip(params, grads):
                n = g.shape[0]
                wgrad(dz[:, r:r + n], x2, g, n, K, M, (prm,), lda=dz.stride(0))
                r += n
        return dx, None, None, None, None


def linear_qkv(x, wq, wk, wv):
    return _LinearQKV.apply(x, wq, wk, wv, (wq, wk, wv))


# ----------------------------------------------------------------------------------------
def _ln_backward(ctx, dy2, x2, mean, rstd, dr):
    """dx = LN'(dy) [+ dres]; with a producer hand-off also that layer's dz / bias gradient (see _producer_of)."""
    weight, bias = ctx.params
    rows, D = x2.shape
    if not getattr(ctx, "mid_block", False):
        flush_wgrads()            # a block's entry LayerNorm is the last node of that block's backward
    dx = torch.empty((rows, D), dtype=torch.float32, device=dy2.device)
    sc, cnt = colsum_scratch(dy2.device)
    prod = getattr(ctx, "producer", None)
    if prod is not None and getattr(prod, "fuse_info", None) is not None:
        _, mask, act, pbias = prod.fuse_info
        dz = torch.empty((rows, D), dtype=BF16, device=dy2.device)
        dbz = grad_of(pbias) if (pbias is not None and pbias.requires_grad) else None
        call("b200_layernorm_bwd_fused", dy2.data_ptr(), x2.data_ptr(), weight.data_ptr(), mean.data_ptr(),
             rstd.data_ptr(), ptr(dr), dx.data_ptr(), dz.data_ptr(), grad_of(weight).data_ptr(),
             grad_of(bias).data_ptr(), rows, D, sc.data_ptr(), cnt.data_ptr(), ptr(mask) if act != 0 else None, act,
             ptr(dbz), 1, stream())
        prod.fused = (dz, dx.data_ptr())
    else:
        call("b200_layernorm_bwd", dy2.data_ptr(), x2.data_ptr(), weight.data_ptr(), mean.data_ptr(),
             rstd.data_ptr(), ptr(dr), dx.data_ptr(), None, grad_of(weight).data_ptr(),
             grad_of(bias).data_ptr(), rows, D, sc.data_ptr(), cnt.data_ptr(), stream())
    _notify(weight, bias)
    return dx


class _LayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps, params, producer=(None, False)):
        D = x.shape[-1]
        x2 = x.reshape(-1, D)
        if x2.dtype != torch.float32:
            x2 = x2.float()
        x2 = x2.contiguous()
        ctx.producer, ctx.mid_block = producer
        rows = x2.shape[0]
        y = torch.empty((rows, D), dtype=BF16, device=x.device)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        call("b200_layernorm_fwd", x2.data_ptr(), weight.data_ptr(), bias.data_ptr(), y.data_ptr(),
             mean.data_ptr(), rstd.data_ptr(), rows, D, float(eps), stream())
        ctx.params = params
        ctx.in_dtype = x.dtype
        ctx.save_for_backward(x2, mean, rstd)
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, mean, rstd = ctx.saved_tensors
        weight, bias = ctx.params
        rows, D = x2.shape
        dy2 = dy.reshape(rows, D)
        if dy2.dtype != BF16:
            dy2 = dy2.to(BF16)
        dy2 = dy2.contiguous()
        dx = _ln_backward(ctx, dy2, x2, mean, rstd, None)
        dx = dx.view(dy.shape)
        if ctx.in_dtype != torch.float32:
            dx = dx.to(ctx.in_dtype)
        return dx, None, None, None, None, None


def layernorm(x, weight, bias, eps):
    return _LayerNorm.apply(x, weight, bias, eps, (weight, bias), _producer_of(x))


class _LayerNormResidual(torch.autograd.Function):
    """``(x, LayerNorm(x))``: the pre-LN block needs both - the normalised activations for the
    sub-layer and x itself for the residual add that the sub-layer's last GEMM performs in its
    epilogue.  Routing both through one autograd node lets the backward kernel add the residual
    branch's gradient while it writes dx (``dres`` of b200_layernorm_bwd) instead of autograd
    launching a separate fp32 add over the whole residual stream."""

    @staticmethod
    def forward(ctx, x, weight, bias, eps, params, producer=(None, False)):
        D = x.shape[-1]
        x2 = x.reshape(-1, D)
        if x2.dtype != torch.float32 or not x2.is_contiguous():
            x2 = x2.float().contiguous()
        ctx.producer, ctx.mid_block = producer
        rows = x2.shape[0]
        y = torch.empty((rows, D), dtype=BF16, device=x.device)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        call("b200_layernorm_fwd", x2.data_ptr(), weight.data_ptr(), bias.data_ptr(), y.data_ptr(),
             mean.data_ptr(), rstd.data_ptr(), rows, D, float(eps), stream())
        ctx.params = params
        ctx.in_dtype = x.dtype
        ctx.save_for_backward(x2, mean, rstd)
        return x.view_as(x), y.view(x.shape)

    @staticmethod
    def backward(ctx, dres, dy):
        x2, mean, rstd = ctx.saved_tensors
        weight, bias = ctx.params
        rows, D = x2.shape
        if dy is None:
            return dres, None, None, None, None, None
        dy2 = dy.reshape(rows, D)
        if dy2.dtype != BF16:
            dy2 = dy2.to(BF16)
        dy2 = dy2.contiguous()
        dr = None
        if dres is not None:
            dr = dres.reshape(rows, D)
            if dr.dtype != torch.float32 or not dr.is_contiguous():
                dr = dr.float().contiguous()
        dx = _ln_backward(ctx, dy2, x2, mean, rstd, dr)
        dx = dx.view(dy.shape)
        if ctx.in_dtype != torch.float32:
            dx = dx.to(ctx.in_dtype)
        return dx, None, None, None, None, None


def layernorm_residual(x, weight, bias, eps):
    """-> (x for the residual branch, LayerNorm(x) as bf16)"""
    return _LayerNormResidual.apply(x, weight, bias, eps, (weight, bias), _producer_of(x))


# ----------------------------------------------------------------------------------------
class _Embedding(torch.autograd.Function):
    @staticmethod
    def forward(ctx, input_ids, position_ids, tok_w, pos_w, params):
        if _wgrad_queue:          # a backward pass that ended without its engine callback (e.g. it raised)
            flush_wgrads()
        ids = input_ids.reshape(-1).contiguous()
        pos = position_ids.reshape(-1).contiguous()
        n, D = ids.numel(), tok_w.shape[1]
        out = torch.empty((n, D), dtype=torch.float32, device=ids.device)
        call("b200_embedding_fwd", ids.data_ptr(), pos.data_ptr(), tok_w.data_ptr(), pos_w.data_ptr(),
             out.data_ptr(), n, D, tok_w.shape[0], pos_w.shape[0], stream())
        ctx.params = params
        ctx.save_for_backward(ids, pos)
        return out.view(*input_ids.shape, D)

    @staticmethod
    def backward(ctx, dout):
        ids, pos = ctx.saved_tensors
        tok_w, pos_w = ctx.params
        D = tok_w.shape[1]
        d2 = dout.reshape(-1, D)
        if d2.dtype != torch.float32:
            d2 = d2.float()
        d2 = d2.contiguous()
        call("b200_embedding_bwd", ids.data_ptr(), pos.data_ptr(), d2.data_ptr(),
             grad_of(tok_w).data_ptr(), grad_of(pos_w).data_ptr(), ids.numel(), D,
             tok_w.shape[0], pos_w.shape[0], stream())
        _notify(tok_w, pos_w)
        return None, None, None, None, None


def embedding(input_ids, position_ids, tok_w, pos_w):
    return _Embedding.apply(input_ids, position_ids, tok_w, pos_w, (tok_w, pos_w))


# ----------------------------------------------------------------------------------------
class _Attention(torch.autograd.Function):
    """Causal flash attention over a packed [N, S, 3*h*d] qkv tensor (csrc/attention.cu)."""

    @staticmethod
    def forward(ctx, qkv, heads, head_dim, scale, key_mask):
        N, S, _ = qkv.shape
        qkv = qkv.contiguous()
        out = torch.empty((N, S, heads * head_dim), dtype=BF16, device=qkv.device)
        lse = torch.empty((N, heads, S), dtype=torch.float32, device=qkv.device)
        km = None
        if key_mask is not None:
            km = key_mask.to(torch.uint8).contiguous()
        call("b200_attn_fwd", qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), N, S, heads, head_dim,
             float(scale), ptr(km), stream())
        ctx.dims = (N, S, heads, head_dim, scale)
        ctx.save_for_backward(qkv, out, lse, km)
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse, km = ctx.saved_tensors
        N, S, heads, head_dim, scale = ctx.dims
        dout = dout.contiguous()
        if dout.dtype != BF16:
            dout = dout.to(BF16)
        dqkv = torch.empty_like(qkv)
        # scratch: delta = rowsum(dO * O) and the fp32 dQ accumulator reduced across key tiles
        delta = torch.empty((N, heads, S), dtype=torch.float32, device=qkv.device)
        dq_acc = torch.empty((N, S, heads * head_dim), dtype=torch.float32, device=qkv.device)
        call("b200_attn_bwd", qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(),
             dqkv.data_ptr(), delta.data_ptr(), dq_acc.data_ptr(), N, S, heads, head_dim, float(scale),
             ptr(km), stream())
        return dqkv, None, None, None, None


def attention(qkv, heads, head_dim, scale, key_mask=None):
    if not _lib.has("b200_attn_fwd"):
        # bring-up only: the library was built without csrc/attention.cu
        N, S, _ = qkv.shape
        q, k, v = (t.reshape(N, S, heads, head_dim).transpose(1, 2) for t in qkv.chunk(3, dim=-1))
        from . import attention_reference
        return attention_reference(q, k, v, scale, key_mask).transpose(1, 2).reshape(N, S, heads * head_dim)
    return _Attention.apply(qkv, heads, head_dim, scale, key_mask)


# ----------------------------------------------------------------------------------------
_LM_FUSED = _os.environ.get("B200GPT_LM_FUSED", "1") != "0"     # 0: the older logits-buffer + in-place CE path
_LM_CHECKS_LEFT = [2]                                             # un-captured backward calls that verify dloss


def lm_state(weight: torch.Tensor) -> torch.Tensor:
    """The 4 floats the LM-head kernels keep with a weight: [max row norm, softmax-shift gap in use, gap
    observed by the last call, rows the last call had to recompute] (csrc/elementwise.cu)."""
    st = getattr(weight, "_b200_lmstate", None)
    if st is None or st.device != weight.device:
        st = torch.zeros(4, dtype=torch.float32, device=weight.device)
        weight._b200_lmstate = st
    return st


class _LMHeadCE(torch.autograd.Function):
    """Fused LM head + cross-entropy (+ argmax accuracy); the logits never exist in memory.

    ``b200_lmhead_gemm_bf16`` computes the logits tile by tile in tensor memory and its epilogue turns
    them straight into ``E = exp(logit - shift[row])`` (bf16) plus per-row (sum, max) statistics and
    the target logit; the shift is fixed before the GEMM from ``|x| * max|w|`` and the gap observed on the
    previous call, so no second pass over a row is needed (rows the estimate misses are recomputed
    exactly by ``b200_lmhead_finish``).  ``finish`` produces loss / accuracy and the per-row factor
    ``grad_scale / (n_valid * sum)``; the gradients are then two GEMMs over E:
    ``dX = diag(rscale) (E' W)`` and ``dW += E'^T (rscale * X)`` with ``E' = E - sum * onehot``.
    E (rows x vocab bf16, one reused chunk of ``B200GPT_LM_CHUNK`` rows) is written once and read
    twice; evaluation stores nothing.  The gradient is produced in the forward pass; backward only
    hands dX over, so the incoming gradient of the loss must equal ``grad_scale`` (1.0 for plain
    training; pipeline engines pass 1/num_microbatches) - verified on the first un-captured calls
    and always under ``B200GPT_DEBUG=1``.
    """

    @staticmethod
    def forward(ctx, x, weight, targets, ignore_index, grad_scale, params, grad_on):
        if not _LM_FUSED:
            return _lmhead_ce_unfused(ctx, x, weight, targets, ignore_index, grad_scale, grad_on)
        x2 = _as_bf16_2d(x)
        M, D = x2.shape
        wb = bf16_of(weight)
        V = wb.shape[0]
        ld = (V + 127) // 128 * 128
        t = targets.reshape(-1).contiguous()
        dev = x.device
        st = stream()
        stats = torch.empty(4, dtype=torch.float32, device=dev)  # n_valid, loss_sum, n_correct
        lmstate = lm_state(weight)
        call("b200_lmhead_begin", t.data_ptr(), stats.data_ptr(), lmstate.data_ptr(), ignore_index, M, st)
        xnorm = torch.empty(M, dtype=torch.float32, device=dev)
        call("b200_lmhead_norms", x2.data_ptr(), x2.stride(0), wb.data_ptr(), wb.stride(0), M, V, D,
             xnorm.data_ptr(), lmstate.data_ptr(), st)
        want_grad = grad_on and (ctx.needs_input_grad[0] or ctx.needs_input_grad[1])
        rows = min(_LM_CHUNK_ROWS, M)
        max_parts = (V + 127) // 128 + 2
        part = torch.empty((max_parts, rows, 2), dtype=torch.float32, device=dev)
        tgt_val = torch.empty(rows, dtype=torch.float32, device=dev)
        E = xs = rscale = dx = None
        gw = None
        if want_grad:
            E = torch.empty((rows, ld), dtype=BF16, device=dev)
            xs = torch.empty((rows, D), dtype=BF16, device=dev)
            rscale = torch.empty(rows, dtype=torch.float32, device=dev)
            dx = torch.empty((M, D), dtype=BF16, device=dev)
            gw = grad_of(weight) if weight.requires_grad else None
        nparts = _ctypes.c_int(0)
        for r0 in range(0, M, rows):
            r = min(rows, M - r0)
            xc = x2[r0:r0 + r]
            call("b200_lmhead_gemm_bf16", xc.data_ptr(), wb.data_ptr(), ptr(E), r, V, D, xc.stride(0), wb.stride(0), ld,
                 xnorm.data_ptr() + 4 * r0, lmstate.data_ptr(), part.data_ptr(), max_parts, _ctypes.addressof(nparts),
                 t.data_ptr() + 8 * r0, tgt_val.data_ptr(), 0, st)
            call("b200_lmhead_finish", part.data_ptr(), nparts.value, r, xnorm.data_ptr() + 4 * r0, tgt_val.data_ptr(),
                 t.data_ptr() + 8 * r0, ignore_index, stats.data_ptr(), lmstate.data_ptr(), ptr(E), ld, V,
                 xc.data_ptr(), xc.stride(0), D, wb.data_ptr(), wb.stride(0), ptr(xs), ptr(rscale),
                 float(grad_scale), int(want_grad), st)
            if want_grad:
                Ec = E[:r]
                if r >= 8192:
                    # enough row tiles to fill the SMs: dX straight to bf16, row scale in the epilogue
                    call("b200_gemm_rowscale_bf16", Ec.data_ptr(), wb.data_ptr(), dx.data_ptr() + 2 * r0 * D, r, D, V,
                         ld, wb.stride(0), D, 1, rscale.data_ptr(), 0, st)
                else:
                    # few output tiles, K = vocab: split-K into fp32 (TMA reduce-add), scale + cast afterwards
                    dx32 = torch.zeros((r, D), dtype=torch.float32, device=dev)
                    gemm(Ec, wb, dx32, r, D, V, b_mn_major=True, accumulate=True)
                    dx[r0:r0 + r] = (dx32 * rscale[:r, None]).to(BF16)
                if gw is not None:
                    gemm(Ec, xs[:r], gw, V, D, r, a_mn_major=True, b_mn_major=True, accumulate=True)   # dW += E'^T . Xs
        if gw is not None:
            _notify(weight)
        loss = stats[1] / stats[0].clamp(min=1.0)
        ctx.save_for_backward(dx)
        ctx.x_shape, ctx.in_dtype, ctx.grad_scale = x.shape, x.dtype, float(grad_scale)
        ctx.mark_non_differentiable(stats)
        return loss, stats

    @staticmethod
    def backward(ctx, dloss, _dstats):
        (dx,) = ctx.saved_tensors
        if dx is None:
            return None, None, None, None, None, None, None
        debug = _os.environ.get("B200GPT_DEBUG") == "1"
        if (debug or _LM_CHECKS_LEFT[0] > 0) and not torch.cuda.is_current_stream_capturing():
            _LM_CHECKS_LEFT[0] -= 1
            got = float(dloss)
            if abs(got - ctx.grad_scale) > 1e-6 * max(1.0, abs(ctx.grad_scale)):
                raise RuntimeError(
                    f"lmhead_cross_entropy: loss.backward() arrived with gradient {got} but the op was told "
                    f"grad_scale={ctx.grad_scale}; its weight gradient is produced in the forward pass, so pass the "
                    "factor you scale the loss by as grad_scale=")
        dx = dx.view(ctx.x_shape)
        if ctx.in_dtype != BF16:
            dx = dx.to(ctx.in_dtype)
        return dx, None, None, None, None, None, None


def _lmhead_ce_unfused(ctx, x, weight, targets, ignore_index, grad_scale, grad_on):
    """B200GPT_LM_FUSED=0: logits chunk -> in-place ``b200_ce_fwd_bwd`` -> dX / dW GEMMs (the pre-fusion path,
    kept as the A/B baseline of the fused kernels; vocabularies up to 57856)."""
    x2 = _as_bf16_2d(x)
    M, D = x2.shape
    wb = bf16_of(weight)
    V = wb.shape[0]
    ld = (V + 127) // 128 * 128
    t = targets.reshape(-1).contiguous()
    dev = x.device
    stats = torch.zeros(4, dtype=torch.float32, device=dev)  # n_valid, loss_sum, n_correct
    call("b200_count_valid", t.data_ptr(), stats.data_ptr(), ignore_index, M, stream())
    want_grad = grad_on and (ctx.needs_input_grad[0] or ctx.needs_input_grad[1])
    rows = min(_LM_CHUNK_ROWS, M)
    logits = torch.empty((rows, ld), dtype=BF16, device=dev)
    one_chunk = M <= rows and M >= 8192
    dx32 = torch.zeros((M, D), dtype=torch.float32, device=dev) if (want_grad and not one_chunk) else None
    dx = torch.empty((M, D), dtype=BF16, device=dev) if (want_grad and one_chunk) else None
    gw = grad_of(weight) if (want_grad and weight.requires_grad) else None
    for r0 in range(0, M, rows):
        r = min(rows, M - r0)
        sub = _LM_SUB_ROWS if _LM_SUB_ROWS > 0 else r
        for s0 in range(0, r, sub):
            sr = min(sub, r - s0)
            gemm(x2[r0 + s0:r0 + s0 + sr], wb, logits[s0:s0 + sr], sr, V, D)
            call("b200_ce_fwd_bwd", logits.data_ptr() + 2 * s0 * ld, t.data_ptr() + 8 * (r0 + s0), stats.data_ptr(),
                 sr, V, ld, ignore_index, int(want_grad), float(grad_scale), stream())
        if want_grad:
            if one_chunk:
                gemm(logits, wb, dx, r, D, V, b_mn_major=True)                             # dX = dL . W
            else:
                gemm(logits[:r], wb, dx32[r0:r0 + r], r, D, V, b_mn_major=True, accumulate=True)
            if gw is not None:
                gemm(logits[:r], x2[r0:r0 + r], gw, V, D, r, a_mn_major=True, b_mn_major=True,
                     accumulate=True)                                       # dW += dL^T . X
    if gw is not None:
        _notify(weight)
    loss = stats[1] / stats[0].clamp(min=1.0)
    if dx32 is not None:
        dx = dx32.to(BF16)
    ctx.save_for_backward(dx)
    ctx.x_shape, ctx.in_dtype, ctx.grad_scale = x.shape, x.dtype, float(grad_scale)
    ctx.mark_non_differentiable(stats)
    return loss, stats


def lmhead_cross_entropy(x, weight, targets, ignore_index=-100, grad_scale: float = 1.0):
    loss, stats = _LMHeadCE.apply(x, weight, targets, ignore_index, grad_scale, (weight,),
                                  torch.is_grad_enabled())
    return loss, stats[0], stats[2]
