"""Reference-compatible ``data`` module (/root/reference/data.py:7-36) plus the synthetic
token stream used when the HuggingFace hub is unreachable."""
from distributed_pytorch_cookbook_b200.data import (PinnedPrefetcher, ShardedSampler,  # noqa: F401
                                                    SyntheticTokenDataset, SyntheticTokenizer,
                                                    get_dataset, get_tokenizer, make_loader,
                                                    transform_dataset)

try:  # the reference re-exports GPT2Tokenizer from here (utils.py:3, main-single.py:13)
    from transformers import GPT2Tokenizer  # noqa: F401
except Exception:  # pragma: no cover
    GPT2Tokenizer = SyntheticTokenizer
