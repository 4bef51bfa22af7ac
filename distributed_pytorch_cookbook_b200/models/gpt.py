"""GPT decoder LM with the reference's public surface and state_dict layout.

Parity target: /root/reference/models/gpt.py:10-231 (FeedForward :10-41, SelfAttention
:44-105, DecoderLayer :108-135, TransformerDecoder :138-167, Embeddings :169-185,
TransformerDecoderLM :187-231).  Class names, constructor signatures, sub-module
attribute names and parameter shapes are identical, so ``state_dict()`` round-trips
with the reference in both directions (109 fp32 tensors at the default config).

What differs is *how* forward runs.  Every module has two execution paths:

* ``torch`` path - plain PyTorch math, numerically the reference's intended
  computation (with its two construction bugs fixed: reference gpt.py:177-178 uses an
  unset ``self.dim`` and :227 an undefined ``x``).  Used on CPU, for fp32
  (``--disable_amp``) runs and as the oracle in tests.
* ``native`` path - taken when the input is on a CUDA device, the in-tree sm_100a
  library is loaded and bf16 autocast is active (the reference's AMP contract,
  main-single.py:88-90).  Forward then runs on the hand-written kernels in ``csrc/``
  through the autograd wrappers in ``ops``: fused LayerNorm -> packed QKV tcgen05 GEMM
  -> causal flash attention -> GEMM(+bias+residual) -> LayerNorm -> GEMM(+bias+ReLU)
  -> GEMM(+bias+ReLU+residual), and a fused LM-head + cross-entropy that never
  materialises the [tokens, vocab] logits (``TransformerDecoderLM.loss``).
"""
from __future__ import annotations

from math import sqrt
from typing import Callable, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops


def _act_name(fn: Callable) -> Optional[str]:
    """Name of an activation the native MLP kernels fuse (ReLU, GELU in its erf and tanh forms); ``None`` for any
    other callable, which then runs through PyTorch (reference gpt.py:14-15 accepts any callable)."""
    import functools
    if fn is F.relu or fn is torch.relu or isinstance(fn, nn.ReLU):
        return "relu"
    if fn is F.gelu:
        return "gelu"
    if isinstance(fn, nn.GELU):
        return "gelu_tanh" if fn.approximate == "tanh" else "gelu"
    if isinstance(fn, functools.partial) and fn.func is F.gelu and not fn.args:
        appr = fn.keywords.get("approximate", "none")
        if set(fn.keywords) <= {"approximate"}:
            return "gelu_tanh" if appr == "tanh" else "gelu"
    return None


class FeedForward(nn.Module):
    """up_proj -> act -> down_proj -> act -> dropout (reference gpt.py:10-41).

    The activation really is applied twice (after the down projection too); default
    ReLU.  The native path folds bias+activation (and the caller's residual) into the
    GEMM epilogues.
    """

    def __init__(self, dim: int, mult: int = 4, activation: Callable = F.relu,
                 out_dim: Optional[int] = None, bias: bool = True, dropout: float = 0.0):
        super().__init__()
        self.dim = dim
        self.mult = mult
        self.activation = activation
        self.bias = bias
        self.out_dim = dim if out_dim is None else out_dim
        self.up_proj = nn.Linear(dim, dim * mult, bias=bias)
        self.down_proj = nn.Linear(dim * mult, self.out_dim, bias=bias)
        self.dropout = nn.Dropout(dropout)

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        act = _act_name(self.activation)
        if ops.use_native(x) and act is not None and self.bias and self.dropout.p == 0.0:
            return ops.mlp(x, self.up_proj.weight, self.up_proj.bias, self.down_proj.weight, self.down_proj.bias,
                           residual=residual, activation=act)
        if ops.use_native(x) and act == "relu" and self.dropout.p == 0.0:
            h = ops.linear(x, self.up_proj.weight, self.up_proj.bias, activation=act)
            return ops.linear(h, self.down_proj.weight, self.down_proj.bias, activation=act,
                              residual=residual, out_fp32=residual is not None)
        h = self.activation(self.up_proj(x))
        h = self.dropout(self.activation(self.down_proj(h)))
        return h if residual is None else h + residual


class SelfAttention(nn.Module):
    """Causal multi-head self-attention (reference gpt.py:44-105).

    Separate bias-free ``to_q/to_k/to_v`` Linear layers (kept separate for the
    state_dict) and a biased ``to_out``.  Causal mask is additive -1e9, the optional
    key-padding mask (True = pad) fills with ``finfo.min`` (reference :83-95).  The
    native path runs one packed QKV GEMM and a flash-attention kernel; the S x S score
    matrix is never written to memory and no mask is built on the host.
    """

    def __init__(self, dim: int, head_dim: int, heads: int, qkv_bias: bool = False,
                 dropout: float = 0.0):
        super().__init__()
        self.dim = dim
        self.head_dim = head_dim
        self.heads = heads
        self.qkv_bias = qkv_bias
        inner = head_dim * heads
        self.to_q = nn.Linear(dim, inner, bias=qkv_bias)
        self.to_k = nn.Linear(dim, inner, bias=qkv_bias)
        self.to_v = nn.Linear(dim, inner, bias=qkv_bias)
        self.to_out = nn.Linear(inner, dim, bias=True)
        self.dropout = nn.Dropout(dropout)
        self.attn_scale = 1 / sqrt(head_dim)

    def forward(self, x: torch.Tensor, mask: Optional[torch.Tensor] = None,
                residual: Optional[torch.Tensor] = None) -> torch.Tensor:
        if ops.use_native(x) and not self.qkv_bias and self.dropout.p == 0.0:
            qkv = ops.linear_qkv(x, self.to_q.weight, self.to_k.weight, self.to_v.weight)
            o = ops.attention(qkv, self.heads, self.head_dim, self.attn_scale, key_mask=mask)
            return ops.linear(o, self.to_out.weight, self.to_out.bias, residual=residual,
                              out_fp32=residual is not None, handoff="block")
        N, S, _ = x.shape
        h, d = self.heads, self.head_dim
        q = self.to_q(x).view(N, S, h, d).transpose(1, 2)
        k = self.to_k(x).view(N, S, h, d).transpose(1, 2)
        v = self.to_v(x).view(N, S, h, d).transpose(1, 2)
        o = ops.attention_reference(q, k, v, self.attn_scale, key_mask=mask)
        o = o.transpose(1, 2).reshape(N, S, h * d)
        o = self.dropout(self.to_out(o))
        return o if residual is None else o + residual


    def forward_cached(self, x: torch.Tensor, past=None):
        """Incremental decoding: attend the ``S`` new positions of ``x`` to the cached keys / values
        plus themselves.  Returns ``(out, (k, v))`` with k, v ``[N, h, L, d]`` covering all ``L``
        positions so far.  Same maths as ``forward`` (reference gpt.py:75-103), without recomputing
        the prefix - the reference's ``generate`` has no cache (utils.py:63-65)."""
        N, S, _ = x.shape
        h, d = self.heads, self.head_dim
        q = self.to_q(x).view(N, S, h, d).transpose(1, 2)
        k = self.to_k(x).view(N, S, h, d).transpose(1, 2)
        v = self.to_v(x).view(N, S, h, d).transpose(1, 2)
        if past is not None:
            k = torch.cat([past[0], k], dim=2)
            v = torch.cat([past[1], v], dim=2)
        L = k.shape[2]
        logits = (q @ k.transpose(-1, -2)) * self.attn_scale
        if S > 1:   # new position i (absolute L - S + i) sees keys 0 .. L - S + i
            causal = torch.ones(S, L, device=x.device, dtype=torch.bool).tril_(diagonal=L - S)
            logits = logits + torch.where(causal, 0.0, -1e9).to(logits.dtype)
        o = F.softmax(logits, dim=-1) @ v
        o = o.transpose(1, 2).reshape(N, S, h * d)
        return self.dropout(self.to_out(o)), (k, v)


class DecoderLayer(nn.Module):
    """Pre-LN block ``x = attn(norm1(x)) + x; x = fc(norm2(x)) + x`` (reference :108-135)."""

    def __init__(self, dim: int, head_dim: int, heads: int, dropout: float = 0.0):
        super().__init__()
        self.dim = dim
        self.head_dim = head_dim
        self.heads = heads
        self.attn = SelfAttention(dim=dim, head_dim=head_dim, heads=heads, dropout=dropout)
        self.norm1 = nn.LayerNorm(dim)
        self.fc = FeedForward(dim=dim, dropout=dropout)
        self.norm2 = nn.LayerNorm(dim)

    def forward(self, x: torch.Tensor, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        if ops.use_native(x):
            x, h = ops.layernorm_residual(x, self.norm1.weight, self.norm1.bias, self.norm1.eps)
            x = self.attn(h, mask, residual=x)
            x, h = ops.layernorm_residual(x, self.norm2.weight, self.norm2.bias, self.norm2.eps)
            return self.fc(h, residual=x)
        x = self.attn(self.norm1(x), mask) + x
        x = self.fc(self.norm2(x)) + x
        return x


    def forward_cached(self, x: torch.Tensor, past=None):
        a, present = self.attn.forward_cached(self.norm1(x), past)
        x = a + x
        x = self.fc(self.norm2(x)) + x
        return x, present


class TransformerDecoder(nn.Module):
    """``num_layers`` DecoderLayers run in sequence (reference gpt.py:138-167)."""

    def __init__(self, dim: int, head_dim: int, heads: int, num_layers: int, dropout: float = 0.0):
        super().__init__()
        self.dim = dim
        self.head_dim = head_dim
        self.heads = heads
        self.num_layers = num_layers
        self.layers = nn.ModuleList(
            [DecoderLayer(dim=dim, head_dim=head_dim, heads=heads, dropout=dropout)
             for _ in range(num_layers)])

    def forward(self, x: torch.Tensor, mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        for layer in self.layers:
            x = layer(x, mask)
        return x


class Embeddings(nn.Module):
    """Token + learned absolute position embeddings, summed (reference gpt.py:169-185)."""

    def __init__(self, dim: int, vocab_size: int, max_position_embeddings: int):
        super().__init__()
        self.dim = dim
        self.input_embeddings = nn.Embedding(vocab_size, dim)
        self.position_embeddings = nn.Embedding(max_position_embeddings, dim)

    def forward(self, input_ids: torch.Tensor, position_ids: torch.Tensor) -> torch.Tensor:
        if ops.use_native(input_ids):
            return ops.embedding(input_ids, position_ids, self.input_embeddings.weight,
                                 self.position_embeddings.weight)
        return self.input_embeddings(input_ids) + self.position_embeddings(position_ids)


class TransformerDecoderLM(nn.Module):
    """embeddings -> decoder -> norm_out -> lm_head (untied) (reference gpt.py:187-231).

    ``forward`` returns logits ``[N, S, vocab]`` like the reference.  ``loss`` is the
    training entry point of this framework: it returns the mean cross-entropy over
    non-ignored targets (plus the valid-token and correct-prediction counts used for
    validation accuracy, reference main-single.py:127-131) and, on the native path,
    fuses the LM head into the loss so the logits are never materialised.
    """

    def __init__(self, dim: int, head_dim: int, heads: int, num_layers: int, vocab_size: int,
                 max_position_embeddings: int, dropout: float = 0.0):
        super().__init__()
        self.dim = dim
        self.head_dim = head_dim
        self.heads = heads
        self.num_layers = num_layers
        self.vocab_size = vocab_size
        self.max_position_embeddings = max_position_embeddings
        self.dropout = dropout
        self.embeddings = Embeddings(dim=dim, vocab_size=vocab_size,
                                     max_position_embeddings=max_position_embeddings)
        self.decoder = TransformerDecoder(num_layers=num_layers, dim=dim, head_dim=head_dim,
                                          heads=heads, dropout=dropout)
        self.norm_out = nn.LayerNorm(dim)
        self.lm_head = nn.Linear(dim, vocab_size, bias=False)

    # -- pieces reused by the pipeline partitioner -------------------------------------
    def hidden(self, input_ids, position_ids, mask=None) -> torch.Tensor:
        x = self.embeddings(input_ids, position_ids)
        return self.decoder(x, mask=mask)

    def head_input(self, x: torch.Tensor) -> torch.Tensor:
        if ops.use_native(x):
            return ops.layernorm(x, self.norm_out.weight, self.norm_out.bias, self.norm_out.eps)
        return self.norm_out(x)

    def forward(self, input_ids: torch.Tensor, position_ids: torch.Tensor,
                mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        x = self.head_input(self.hidden(input_ids, position_ids, mask))
        if ops.use_native(x):
            return ops.linear(x, self.lm_head.weight, None)
        return self.lm_head(x)

    @torch.no_grad()
    def forward_cached(self, input_ids: torch.Tensor, position_ids: torch.Tensor, past=None):
        """Logits for the new positions only, given the per-layer key/value cache ``past`` returned
        by the previous call (``None`` for the prompt).  -> ``(logits [N, S_new, vocab], past)``."""
        with ops.force_reference():      # single-token steps: plain PyTorch maths, any device / dtype
            x = self.embeddings(input_ids, position_ids)
            present = []
            for i, layer in enumerate(self.decoder.layers):
                x, kv = layer.forward_cached(x, None if past is None else past[i])
                present.append(kv)
            return self.lm_head(self.norm_out(x)), present

    def loss(self, input_ids: torch.Tensor, position_ids: torch.Tensor, targets: torch.Tensor,
             mask: Optional[torch.Tensor] = None, ignore_index: int = -100):
        """Mean CE over non-ignored targets -> ``(loss, n_valid, n_correct)``."""
        x = self.head_input(self.hidden(input_ids, position_ids, mask))
        return ops.lmhead_cross_entropy(x, self.lm_head.weight, targets, ignore_index=ignore_index)
