from .hf import get_dataset, get_tokenizer, transform_dataset  # noqa: F401
from .sampler import ShardedSampler  # noqa: F401
from .synthetic import (PinnedPrefetcher, SyntheticTokenDataset, SyntheticTokenizer,  # noqa: F401
                        make_loader)
