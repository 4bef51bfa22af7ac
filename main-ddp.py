#!/usr/bin/env python
"""
Data-parallel recipe - same flags as the reference's main-ddp.py:191-203.  Run with torchrun:

    torchrun --nnodes=1 --nproc_per_node=8 --master-addr 127.0.0.1 main-ddp.py --synthetic

One process per GPU; gradients are all-reduced bucket by bucket with the framework's own
NVLink kernels (``--comm nccl`` selects NCCL instead); on a CPU box it runs over gloo.
"""
from types import SimpleNamespace

from distributed_pytorch_cookbook_b200.parallel.ddp import DDPStrategy
from distributed_pytorch_cookbook_b200.trainer import Trainer
from distributed_pytorch_cookbook_b200.utils import build_parser, cleanup_mp, init_mp, world_info


def main(args: SimpleNamespace):
    rank, device_id = init_mp()
    _, _, world = world_info()
    try:
        return Trainer(args, DDPStrategy(), rank=rank, world=world, device=device_id, data_parallel=True).run()
    finally:
        cleanup_mp()


if __name__ == "__main__":
    main(build_parser("ddp").parse_args())
