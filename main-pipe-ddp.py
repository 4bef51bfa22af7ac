#!/usr/bin/env python
"""
Pipeline x data parallel recipe (2-D mesh) - the reference's main-pipe-ddp.py is an empty stub
(main-pipe-ddp.py:1); the intent is pipeline stages replicated data-parallel.

    torchrun --nnodes=1 --nproc_per_node=8 --master-addr 127.0.0.1 main-pipe-ddp.py --synthetic

Default mesh: 2 stages x (world/2) replicas; rank = stage * dp + replica.  Stage gradients are
all-reduced over the replica group with the DDP bucket reducer (own NVLink kernels or NCCL).
"""
from types import SimpleNamespace

from distributed_pytorch_cookbook_b200.parallel.pipe import PipeStrategy
from distributed_pytorch_cookbook_b200.trainer import Trainer
from distributed_pytorch_cookbook_b200.utils import build_parser, cleanup_mp, init_mp, world_info


def main(args: SimpleNamespace):
    rank, device_id = init_mp()
    _, _, world = world_info()
    stages = args.num_stages or 2
    assert world % stages == 0, f"world size {world} must be a multiple of the stage count {stages}"
    dp = world // stages
    try:
        strategy = PipeStrategy(num_stages=stages, num_microbatches=args.num_microbatches,
                                schedule=args.schedule, dp_size=dp)
        # data is sharded over the dp dimension only
        return Trainer(args, strategy, rank=rank % dp, world=dp, device=device_id, data_parallel=dp > 1,
                       is_main=rank == 0).run()
    finally:
        cleanup_mp()


if __name__ == "__main__":
    main(build_parser("pipe-ddp").parse_args())
