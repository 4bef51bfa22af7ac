"""TinyStories + GPT-2 BPE loaders with the reference's signatures.

Parity: /root/reference/data.py:7-14 (``get_dataset``), :18-20 (``get_tokenizer``),
:23-36 (``transform_dataset``).  They need the HuggingFace hub; with no network they
raise a clear error telling the caller to pass ``--synthetic`` (``data.synthetic``).
Imports of ``datasets``/``transformers`` are lazy so the rest of the framework works
where those packages (or the network) are missing.
"""
from __future__ import annotations

from typing import Optional, Union


def _offline_error(what: str, e: Exception):
    raise RuntimeError(
        f"{what} needs the HuggingFace hub (or a local cache) and it is unreachable: {e!r}. "
        "Use the synthetic token stream instead (--synthetic).") from e


def get_dataset(name: str = "roneneldan/TinyStories", slice_size: Optional[Union[str, int]] = None):
    try:
        import datasets
        train = datasets.load_dataset(name, split=f"train[:{slice_size}]" if slice_size is not None else "train")
        val = datasets.load_dataset(name, split="validation")
    except Exception as e:  # network / missing package
        _offline_error("get_dataset", e)
    return train, val


def get_tokenizer(name: str = "roneneldan/TinyStories-1M", max_length: int = 512):
    try:
        from transformers import GPT2Tokenizer
        return GPT2Tokenizer.from_pretrained(name, model_max_length=max_length)
    except Exception as e:
        _offline_error("get_tokenizer", e)


def transform_dataset(dataset, tokenizer, max_length: int = 512, num_proc: int = 8):
    def _tokenize(example):
        return tokenizer(example["text"], padding="max_length", max_length=max_length, truncation=True)

    dataset = dataset.map(_tokenize, batched=True, remove_columns=["text"], num_proc=num_proc)
    dataset.set_format("pt")
    return dataset
