"""Process bootstrap, seeding and teardown.

The reference duplicates ``init_mp``/``cleanup_mp`` in every recipe (main-ddp.py:25-35,
main-fsdp.py:29-39), maps rank -> ``cuda:{rank % ngpu}`` without ``set_device`` and never
seeds.  Here: torchrun-compatible env:// init, device from ``LOCAL_RANK``, NCCL on GPU
boxes / gloo on CPU, deterministic seeding, idempotent teardown.
"""
from __future__ import annotations

import os
import random
from typing import Tuple

import numpy as np
import torch
import torch.distributed as dist


def world_info() -> Tuple[int, int, int]:
    """(rank, local_rank, world_size) from the torchrun environment (1-process defaults)."""
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)),
            int(os.environ.get("WORLD_SIZE", 1)))


def init_mp(backend: str = None, verbose: bool = True):
    """Initialise the default process group; returns ``(rank, device_id)`` like the
    reference's ``init_mp`` (main-ddp.py:25-31).  Works for world_size == 1 without a
    launcher (falls back to a local TCP store on 127.0.0.1)."""
    rank, local_rank, world = world_info()
    use_cuda = torch.cuda.is_available()
    if backend is None:
        backend = "nccl" if use_cuda else "gloo"
    if use_cuda:
        device_id = f"cuda:{local_rank % torch.cuda.device_count()}"
        torch.cuda.set_device(device_id)
    else:
        device_id = "cpu"
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        kwargs = {}
        if use_cuda and backend == "nccl":
            kwargs["device_id"] = torch.device(device_id)
        dist.init_process_group(backend, rank=rank, world_size=world, **kwargs)
    if verbose:
        print(f"Worker {rank} starting up. Using device_id '{device_id}'.", flush=True)
    return rank, device_id


def cleanup_mp():
    if dist.is_initialized():
        dist.destroy_process_group()


def is_main() -> bool:
    return (not dist.is_initialized()) or dist.get_rank() == 0


def seed_everything(seed: int = 0, rank_offset: int = 0):
    """Seed python / numpy / torch.  Model init uses the same seed on every rank (so DDP
    replicas start identical without a broadcast); data order adds ``rank_offset``."""
    random.seed(seed + rank_offset)
    np.random.seed((seed + rank_offset) % (2 ** 32))
    torch.manual_seed(seed + rank_offset)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed + rank_offset)
