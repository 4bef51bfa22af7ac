"""Canonical checkpoints.

The reference saves ``torch.save(model.state_dict())`` once at the end of training
(main-single.py:147-151, main-ddp.py:179-185, main-fsdp.py:194-200) and, because it
wraps the model first, its files carry ``_orig_mod.`` / ``module.`` prefixes.  We always
save the un-prefixed fp32 key set of ``TransformerDecoderLM`` so that
``TransformerDecoderLM(...).load_state_dict(torch.load(path))`` works in the reference,
and our loader strips the wrapper prefixes so reference-produced files load here.
Resume state (optimizer moments, step counter, epochs done) is additive: the Trainer writes it beside the
weights as ``<checkpoint>.resume.rank<r>`` (one per rank - FSDP shards and pipeline stages own different
slices), so the weights file stays reference-loadable; ``load_checkpoint`` hands the rank's file back and
``--resume`` restores it when the world layout matches (weights only otherwise).
"""
from __future__ import annotations

from collections import OrderedDict
from datetime import datetime
from pathlib import Path
from typing import Dict, Optional

import torch

_PREFIXES = ("_orig_mod.", "module.", "_fsdp_wrapped_module.", "_checkpoint_wrapped_module.")


def strip_prefixes(state: Dict[str, torch.Tensor]) -> "OrderedDict[str, torch.Tensor]":
    out = OrderedDict()
    for k, v in state.items():
        changed = True
        while changed:
            changed = False
            for p in _PREFIXES:
                if k.startswith(p):
                    k = k[len(p):]
                    changed = True
        k = k.replace("._fsdp_wrapped_module.", ".").replace("._orig_mod.", ".")
        out[k] = v
    return out


def canonical_state_dict(model) -> "OrderedDict[str, torch.Tensor]":
    """fp32, CPU, contiguous, un-prefixed copy of the model weights."""
    m = getattr(model, "module", model)
    sd = strip_prefixes(m.state_dict())
    return OrderedDict((k, v.detach().to("cpu", torch.float32).contiguous().clone()) for k, v in sd.items())


def checkpoint_name(now: Optional[datetime] = None) -> str:
    return "checkpoint-" + (now or datetime.now()).strftime("%Y-%m-%d_%H-%M-%S") + ".pt"


def save_checkpoint(state_or_model, directory="checkpoints", name: Optional[str] = None,
                    extra: Optional[dict] = None) -> Path:
    directory = Path(directory)
    directory.mkdir(exist_ok=True, parents=True)
    state = state_or_model if isinstance(state_or_model, dict) else canonical_state_dict(state_or_model)
    path = directory / (name or checkpoint_name())
    torch.save(state, path)
    if extra is not None:
        torch.save(extra, path.with_suffix(".resume.pt"))
    return path


def load_checkpoint(model, path, strict: bool = True, map_location="cpu", rank: int = 0):
    """Load a checkpoint written by us or by any of the reference's recipes.  -> (load result, resume state of
    ``rank`` or None)."""
    state = torch.load(path, map_location=map_location, weights_only=True)
    state = strip_prefixes(state)
    m = getattr(model, "module", model)
    with torch.no_grad():
        result = m.load_state_dict(state, strict=strict)
    # parameters changed behind the kernels' back: drop cached bf16 copies
    for p in m.parameters():
        if hasattr(p, "_b200_bf16_version"):
            p._b200_bf16_version = -1
    resume = Path(str(path) + f".resume.rank{rank}")
    if not resume.exists():
        resume = Path(path).with_suffix(".resume.pt")
    extra = torch.load(resume, map_location=map_location, weights_only=False) if resume.exists() else None
    return result, extra
