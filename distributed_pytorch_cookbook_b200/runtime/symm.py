"""Symmetric (peer-mapped) memory over NVLink 5 / NVSwitch.

Allocation and handle exchange go through ``torch.distributed._symmetric_memory`` (CUDA VMM
driver API + the process-group store - no NCCL on the data path); what we keep
are raw addresses: the device-resident table of every peer's base pointer, the NVLS multicast
mapping when the fabric offers one, and a small symmetric flag buffer used by our kernels'
self-resetting cross-GPU barriers (csrc/comm.cu).  These go straight into kernel arguments.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

FLAG_CHANNELS = 256
FLAG_WORDS_PER_CHANNEL = 16   # == MAX_WORLD in csrc/comm.cu

# channel ranges of kernels that may run concurrently
CH_ALLREDUCE = 0      # [0, 64)
CH_RS_ADAMW = 64      # [64, 128)
CH_ALLGATHER = 128    # [128, 192)
CH_BARRIER = 192      # [192, 256)


def symmetric_memory_available(device) -> bool:
    if torch.device(device).type != "cuda" or not dist.is_initialized():
        return False
    try:
        import importlib
        importlib.import_module("torch.distributed._symmetric_memory")
    except Exception:
        return False
    return True


class SymmetricBuffer:
    """One symmetric allocation, rendezvoused over ``group`` (collective)."""

    def __init__(self, numel: int, dtype: torch.dtype, device, group=None, zero: bool = True):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.tensor = symm_mem.empty(numel, dtype=dtype, device=device)
        if zero:
            self.tensor.zero_()
        self.handle = symm_mem.rendezvous(self.tensor, self.group)
        self.rank: int = self.handle.rank
        self.world: int = self.handle.world_size
        self.ptrs_dev: int = int(self.handle.buffer_ptrs_dev)       # device array [world] of base pointers
        self.peer_ptrs = [int(p) for p in self.handle.buffer_ptrs]
        mc = 0
        try:
            if self.handle.has_multicast_support:
                mc = int(self.handle.multicast_ptr)
        except Exception:
            mc = 0
        self.mc_ptr: int = mc

    def peer_tensor(self, peer: int, shape, dtype) -> torch.Tensor:
        return self.handle.get_buffer(peer, shape, dtype)


class SymmContext:
    """Per-group context: the flag buffer shared by all our collective kernels."""

    def __init__(self, device, group=None):
        self.device = torch.device(device)
        self.group = group if group is not None else dist.group.WORLD
        self.flags = SymmetricBuffer(FLAG_CHANNELS * FLAG_WORDS_PER_CHANNEL, torch.int32, device, self.group)
        self.rank, self.world = self.flags.rank, self.flags.world
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def alloc(self, numel: int, dtype: torch.dtype, zero: bool = True) -> SymmetricBuffer:
        buf = SymmetricBuffer(numel, dtype, self.device, self.group, zero=zero)
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)
        return buf

    def barrier(self, channel: int = CH_BARRIER):
        """Device-side barrier across the group on the current stream (no host sync)."""
        from ..ops import _lib
        _lib.call("b200_barrier", self.flags.ptrs_dev, self.rank, self.world, channel, _lib.stream())
