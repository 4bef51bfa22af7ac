"""T0 oracle: the reference's own model source, loaded at test time from the read-only
mount (or the git-ignored baseline/_ref copy) with the two one-token repairs R1/R2 of
SURVEY 2.8 applied in memory.  Nothing from the reference is stored in this repo."""
import sys
import types
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CANDIDATES = [Path("/root/reference"), ROOT / "baseline" / "_ref"]


def reference_root():
    for c in CANDIDATES:
        if (c / "models" / "gpt.py").exists():
            return c
    return None


def load_reference_gpt():
    root = reference_root()
    if root is None:
        return None
    src = (root / "models" / "gpt.py").read_text()
    src = src.replace("nn.Embedding(vocab_size, self.dim)", "nn.Embedding(vocab_size, dim)")           # R1
    src = src.replace("nn.Embedding(max_position_embeddings, self.dim)", "nn.Embedding(max_position_embeddings, dim)")
    src = src.replace("x = self.embeddings(x, position_ids)", "x = self.embeddings(input_ids, position_ids)")  # R2
    mod = types.ModuleType("_reference_gpt")
    exec(compile(src, str(root / "models" / "gpt.py"), "exec"), mod.__dict__)
    return mod


def load_reference_utils():
    root = reference_root()
    if root is None:
        return None
    src = (root / "utils.py").read_text()
    src = src.replace("from models import TransformerDecoderLM", "TransformerDecoderLM = object")
    src = src.replace("from data import GPT2Tokenizer", "GPT2Tokenizer = object")
    mod = types.ModuleType("_reference_utils")
    exec(compile(src, str(root / "utils.py"), "exec"), mod.__dict__)
    return mod
