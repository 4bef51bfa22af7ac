"""Reference-compatible ``utils`` module: ``prepare_batch`` / ``generate`` with the
reference's signatures (/root/reference/utils.py:5,43) plus the launch / seed / logging /
checkpoint helpers BASELINE.json asks for (the reference duplicates ``init_mp`` /
``cleanup_mp`` in every recipe, main-ddp.py:25-35)."""
from distributed_pytorch_cookbook_b200.utils import (JsonlLogger, StepTimer, build_parser,  # noqa: F401
                                                     canonical_state_dict, cleanup_mp, generate,
                                                     get_logger, init_mp, is_main, load_checkpoint,
                                                     prepare_batch, prepare_batch_device,
                                                     save_checkpoint, seed_everything,
                                                     strip_prefixes, world_info)
