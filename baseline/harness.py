"""Reference-arm loader.

``pip install /root/reference`` is impossible (the reference ships neither setup.py nor
pyproject.toml - outcome recorded in DESIGN.md), so ``baseline/_ref`` (git-ignored, shipped
to the GPU box by gpurun) holds a verbatim copy of the reference's Python files, made by
``ensure_reference_copy()`` whenever /root/reference is mounted.  The files are never edited:
the two one-token repairs without which the model cannot even be constructed (R1: unset
``self.dim`` at models/gpt.py:177-178; R2: undefined ``x`` at :227) are applied to the source
text in memory at import time, the synthetic token batches stand in for the HuggingFace
dataset (R8) and the device comes from LOCAL_RANK with a fixed seed (R9).  Everything else -
model code, ``prepare_batch``, the autocast + GradScaler + AdamW step, DDP, torch.compile - is
the reference's own stock path.
"""
from __future__ import annotations

import shutil
import types
from pathlib import Path
from types import SimpleNamespace
from typing import Optional

HERE = Path(__file__).resolve().parent
REF_DIR = HERE / "_ref"
SRC = Path("/root/reference")
FILES = ["models/__init__.py", "models/gpt.py", "utils.py", "data.py", "main-single.py", "main-ddp.py",
         "main-fsdp.py", "main-pipe.py", "main-pipe-ddp.py", "LICENSE", "README.md", "requirements.txt"]


def ensure_reference_copy() -> bool:
    """Copy the reference verbatim into baseline/_ref when the read-only mount is present."""
    if (REF_DIR / "models" / "gpt.py").exists():
        return True
    if not (SRC / "models" / "gpt.py").exists():
        return False
    for f in FILES:
        s = SRC / f
        if s.exists():
            d = REF_DIR / f
            d.parent.mkdir(parents=True, exist_ok=True)
            shutil.copyfile(s, d)
    return True


def reference_unavailable_reason() -> Optional[str]:
    if not ensure_reference_copy():
        return ("baseline/_ref is empty and /root/reference is not mounted "
                "(pip install of the reference is impossible: no setup.py/pyproject.toml)")
    return None


def _exec(path: Path, name: str, replacements) -> types.ModuleType:
    src = path.read_text()
    for a, b in replacements:
        assert a in src, f"repair anchor not found in {path}: {a!r}"
        src = src.replace(a, b)
    mod = types.ModuleType(name)
    mod.__file__ = str(path)
    exec(compile(src, str(path), "exec"), mod.__dict__)
    return mod


def load_reference() -> SimpleNamespace:
    assert ensure_reference_copy(), reference_unavailable_reason()
    gpt = _exec(REF_DIR / "models" / "gpt.py", "_ref_gpt", [
        ("nn.Embedding(vocab_size, self.dim)", "nn.Embedding(vocab_size, dim)"),                          # R1
        ("nn.Embedding(max_position_embeddings, self.dim)", "nn.Embedding(max_position_embeddings, dim)"),  # R1
        ("x = self.embeddings(x, position_ids)", "x = self.embeddings(input_ids, position_ids)"),       # R2
    ])
    utils = _exec(REF_DIR / "utils.py", "_ref_utils", [
        ("from models import TransformerDecoderLM", "TransformerDecoderLM = object"),   # import plumbing only
        ("from data import GPT2Tokenizer", "GPT2Tokenizer = object"),
    ])
    return SimpleNamespace(gpt=gpt, utils=utils,
                           repairs=["R1 gpt.py:177-178 self.dim->dim", "R2 gpt.py:227 x->input_ids",
                                    "R8 synthetic tokens", "R9 LOCAL_RANK device + seed"])
