from .fused_adamw import FusedAdamW, build_optimizer  # noqa: F401
