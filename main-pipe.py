#!/usr/bin/env python
"""
Pipeline-parallel recipe (GPipe, optional 1F1B) - same flags as the reference's
main-pipe.py:224-236.  One process per stage:

    torchrun --nnodes=1 --nproc_per_node=4 --master-addr 127.0.0.1 main-pipe.py --synthetic

Stages = world size, micro-batches = stages (reference ``chunks=num_stages``) unless
``--num_microbatches`` says otherwise; activations travel over NCCL p2p (gloo on CPU).
"""
from types import SimpleNamespace

from distributed_pytorch_cookbook_b200.parallel.pipe import PipeStrategy
from distributed_pytorch_cookbook_b200.trainer import Trainer
from distributed_pytorch_cookbook_b200.utils import build_parser, cleanup_mp, init_mp


def main(args: SimpleNamespace):
    rank, device_id = init_mp()
    try:
        strategy = PipeStrategy(num_stages=args.num_stages, num_microbatches=args.num_microbatches,
                                schedule=args.schedule, dp_size=1)
        # every rank iterates the same data (data_parallel=False): stage 0 uses the inputs,
        # the last stage the targets
        return Trainer(args, strategy, rank=rank, world=1, device=device_id, data_parallel=False).run()
    finally:
        cleanup_mp()


if __name__ == "__main__":
    main(build_parser("pipe").parse_args())
