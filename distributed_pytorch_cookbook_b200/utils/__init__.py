from .batch import generate, prepare_batch, prepare_batch_device  # noqa: F401
from .checkpoint import (canonical_state_dict, load_checkpoint, save_checkpoint,  # noqa: F401
                         strip_prefixes)
from .launch import cleanup_mp, init_mp, is_main, seed_everything, world_info  # noqa: F401
from .logging import JsonlLogger, StepTimer, get_logger  # noqa: F401
from .cli import build_parser  # noqa: F401
