#!/usr/bin/env python
"""
Fully-sharded data-parallel recipe - same flags as the reference's main-fsdp.py:206-219
(including ``--cpu_offload``).  Run with torchrun:

    torchrun --nnodes=1 --nproc_per_node=8 --master-addr 127.0.0.1 main-fsdp.py --synthetic \
        --dim 2048 --head_dim 128 --heads 16 --num_layers 24 --sequence_length 1024 --batch_size 8
"""
from types import SimpleNamespace

from distributed_pytorch_cookbook_b200.parallel.fsdp import FSDPStrategy
from distributed_pytorch_cookbook_b200.trainer import Trainer
from distributed_pytorch_cookbook_b200.utils import build_parser, cleanup_mp, init_mp, world_info


def main(args: SimpleNamespace):
    rank, device_id = init_mp()
    _, _, world = world_info()
    try:
        return Trainer(args, FSDPStrategy(cpu_offload=args.cpu_offload), rank=rank, world=world,
                       device=device_id, data_parallel=True).run()
    finally:
        cleanup_mp()


if __name__ == "__main__":
    main(build_parser("fsdp").parse_args())
