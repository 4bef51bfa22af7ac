#!/usr/bin/env python
"""Single-device recipe (CUDA if available else CPU) - same flags as the reference's
main-single.py:155-167; the loop itself is the shared ``Trainer``.

    python main-single.py --synthetic --epochs 1 --max_steps 20
"""
from types import SimpleNamespace

import torch

from distributed_pytorch_cookbook_b200.trainer import SingleStrategy, Trainer
from distributed_pytorch_cookbook_b200.utils import build_parser


def main(args: SimpleNamespace):
    device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
    return Trainer(args, SingleStrategy(), rank=0, world=1, device=device).run()


if __name__ == "__main__":
    main(build_parser("single").parse_args())
