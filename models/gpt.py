"""Reference-compatible module path ``models.gpt`` (/root/reference/models/gpt.py)."""
from distributed_pytorch_cookbook_b200.models.gpt import *  # noqa: F401,F403
from distributed_pytorch_cookbook_b200.models.gpt import (DecoderLayer, Embeddings, FeedForward,  # noqa: F401
                                                          SelfAttention, TransformerDecoder,
                                                          TransformerDecoderLM)
