"""Batch preparation and greedy generation.

Parity: /root/reference/utils.py:5-39 (``prepare_batch``) and :42-91 (``generate``).
Same signatures and results; ``prepare_batch_device`` is the additive device-side
variant for token tensors that already live on the GPU (synthetic stream, prefetcher).
"""
from __future__ import annotations

from typing import Optional

import torch


def prepare_batch(batch, pad_id: int, device):
    """HF-style batch ``{input_ids, attention_mask}`` -> ``(model kwargs, targets)``.

    Shift-by-one LM batch: inputs ``ids[:, :-1]``, targets ``ids[:, 1:]`` with ``pad_id``
    replaced by -100 (CrossEntropy ``ignore_index``), ``position_ids = arange(S-1)``,
    ``mask = ~attention_mask[:, :-1]`` (True = padding).  Reference utils.py:16-38.
    Source tensors in pinned memory make the ``non_blocking`` copies truly asynchronous.
    """
    input_ids = batch["input_ids"]
    attention_mask = batch["attention_mask"][:, :-1]
    sequence_length = input_ids.shape[1]
    input_ids, targets = input_ids[:, :-1].clone(), input_ids[:, 1:].clone()
    targets[targets == pad_id] = -100
    position_ids = torch.arange(sequence_length - 1).unsqueeze(0).repeat(input_ids.shape[0], 1)
    out = dict(
        input_ids=input_ids.to(device, non_blocking=True),
        position_ids=position_ids.to(device, non_blocking=True),
        mask=~attention_mask.to(dtype=torch.bool, device=device, non_blocking=True),
    )
    return out, targets.to(device, non_blocking=True)


def prepare_batch_device(input_ids: torch.Tensor, pad_id: int,
                         attention_mask: Optional[torch.Tensor] = None):
    """Same result as ``prepare_batch`` for an ``input_ids`` tensor already on its device.

    One H2D copy of the raw ``[B, S]`` ids is all a step needs; the shift, the -100
    fill and the position ids are produced on the device.
    """
    S = input_ids.shape[1]
    inp = input_ids[:, :-1].contiguous()
    tgt = input_ids[:, 1:].clone()
    tgt[tgt == pad_id] = -100
    pos = torch.arange(S - 1, device=input_ids.device).unsqueeze(0).expand(inp.shape[0], -1).contiguous()
    mask = None if attention_mask is None else ~attention_mask[:, :-1].to(torch.bool)
    return dict(input_ids=inp, position_ids=pos, mask=mask), tgt


@torch.inference_mode()
def generate(model, prompt: str, tokenizer, device, max_new_tokens: int = 20, use_cache: bool = False):
    """Greedy (argmax) decoding, reference utils.py:42-91: stops at EOS, prompt truncated to 256
    tokens.  Like the reference it re-runs the whole prefix each step by default; ``use_cache=True``
    keeps per-layer keys / values (``TransformerDecoderLM.forward_cached``) and feeds one token per
    step - same tokens, O(L) instead of O(L^2) work per step."""
    batch = tokenizer([prompt], truncation=True, max_length=256, return_tensors="pt")
    input_ids = batch["input_ids"].to(device)
    sequence_length = input_ids.shape[1]
    position_ids = torch.arange(sequence_length, device=device).unsqueeze(0)
    max_pos = getattr(model, "max_position_embeddings", None)
    cached = use_cache and hasattr(model, "forward_cached")
    past = None
    for _ in range(max_new_tokens):
        if max_pos is not None and sequence_length >= max_pos:
            break
        if cached:
            new = slice(0, None) if past is None else slice(-1, None)
            logits, past = model.forward_cached(input_ids[:, new], position_ids[:, new], past)
        else:
            logits = model(input_ids=input_ids, position_ids=position_ids)
        new_token = logits[0, -1].argmax(dim=-1)
        if int(new_token) == tokenizer.eos_token_id:
            break
        input_ids = torch.cat([input_ids, new_token.view(1, 1).to(input_ids.dtype)], dim=1)
        position_ids = torch.cat(
            [position_ids, torch.full((1, 1), sequence_length, dtype=position_ids.dtype, device=device)], dim=1)
        sequence_length += 1
    return tokenizer.decode(input_ids[0], skip_special_tokens=True)
