from .gpt import (DecoderLayer, Embeddings, FeedForward, SelfAttention,  # noqa: F401
                  TransformerDecoder, TransformerDecoderLM)
