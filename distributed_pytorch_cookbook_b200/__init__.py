"""b200-distributed-gpt: a Blackwell-native distributed GPT training framework.

Capabilities follow vvvm23/distributed-pytorch-cookbook (single / DDP / FSDP / pipeline /
pipeline x DDP recipes around one GPT decoder LM); the implementation is B200-first:
hand-written sm_100a kernels behind a C ABI (``csrc/``, ``include/b200gpt.h``), CUDA
graphs instead of a tracing compiler, flat parameter/gradient arenas in NVLink-mapped
symmetric memory, and collectives fused into the kernels that produce or consume them.
"""
__version__ = "0.1.0"

from . import ops  # noqa: F401
from .models import TransformerDecoderLM  # noqa: F401
