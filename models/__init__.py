"""Reference-compatible import path (``from models import TransformerDecoderLM``,
/root/reference/models/__init__.py:1); implementation lives in the package."""
from distributed_pytorch_cookbook_b200.models import (DecoderLayer, Embeddings, FeedForward,  # noqa: F401
                                                      SelfAttention, TransformerDecoder,
                                                      TransformerDecoderLM)
