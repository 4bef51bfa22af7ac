"""Data-parallel engine: bucketed gradient all-reduce over the flat gradient arena.

Reference: ``DistributedDataParallel(model, device_ids=[dev])`` (main-ddp.py:55) - the c10d
reducer copies grads into 25 MiB buckets, pre-divides, calls ncclAllReduce and copies back; at
the default config that forms two buckets of which the second (73 MiB, closed by the input
embedding) is fully exposed.

Here the gradient arena itself lives in NVLink-mapped symmetric memory and the kernels write
weight gradients straight into it, so a bucket is just ``arena.grad[lo:hi]``: no copy-in, no
copy-out.  Buckets are small (one per decoder layer by default), become ready the moment the
producing wgrad kernel is *queued* (``ops.native.grad_ready_hooks``; autograd post-accumulate
hooks on the PyTorch fallback path) and are reduced in place by ``b200_allreduce`` on a side
stream: NVLS ``multimem.ld_reduce`` + 1/world scale + ``multimem.st`` in one kernel, no NCCL.
The LM-head gradient is complete right after the fused LM-head/CE *forward*, so its 49 MiB
all-reduce overlaps the whole decoder backward.  Transport fallbacks: NCCL (``--comm nccl``)
and gloo on CPU - same bucket plan, ``dist.all_reduce`` per bucket.
"""
from __future__ import annotations

import os
from contextlib import nullcontext
from typing import List, Optional

import torch
import torch.distributed as dist

from ..ops import _lib, native
from ..optim import FusedAdamW
from ..runtime.arena import ParamArena
from ..runtime.symm import CH_ALLREDUCE, SymmContext, symmetric_memory_available


def plan_buckets(params: List[torch.nn.Parameter], offsets: List[int], total: int, cap_elems: int):
    """Contiguous arena slices, built in reverse registration order (~ backward order).

    Returns a list of ``(lo, hi, [param indices])``; bucket 0 is the first to become ready.
    A parameter larger than the cap gets a bucket of its own.
    """
    buckets = []
    hi = total
    cur: List[int] = []
    cur_lo = total
    for i in range(len(params) - 1, -1, -1):
        lo = offsets[i]
        if cur and (hi - lo) > cap_elems:
            buckets.append((cur_lo, hi, cur))
            hi, cur = cur_lo, []
        cur.append(i)
        cur_lo = lo
    if cur:
        buckets.append((cur_lo, hi, cur))
    # make every slice 16-byte granular for the vector kernels
    assert all(lo % 4 == 0 and hi % 4 == 0 for lo, hi, _ in buckets)
    return buckets


class GradReducer:
    def __init__(self, arena: ParamArena, group=None, bucket_mb: float = 4.0, transport: str = "nccl",
                 symm: Optional[SymmContext] = None, grad_buf=None, nblocks: int = 16, comm_dtype: str = "fp32"):
        self.arena, self.group, self.transport = arena, group, transport
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.symm, self.grad_buf, self.nblocks = symm, grad_buf, nblocks
        self.tail_blocks = int(os.environ.get("B200GPT_DDP_TAIL_BLOCKS", "64"))
        cap = max(int(bucket_mb * (1 << 20) / 4), 1)
        self.cap = cap
        # --comm_dtype bf16: gradients cross NVLink as bf16 (fp32 accumulation in the switch) through a symmetric
        # bf16 staging buffer; needs the NVLS multicast mapping, otherwise the fp32 kernel is used
        self.stage_buf = None
        if comm_dtype == "bf16" and transport == "symm" and self.world > 1:
            buf = symm.alloc(arena.total, torch.bfloat16)
            if buf.mc_ptr:
                self.stage_buf = buf
        self.comm_dtype = "bf16" if self.stage_buf is not None else "fp32"
        self.buckets = plan_buckets(arena.params, arena.offsets, arena.total, cap)
        self.bucket_of = {}
        for b, (_, _, idxs) in enumerate(self.buckets):
            for i in idxs:
                self.bucket_of[id(arena.params[i])] = b
        self.remaining = [len(idxs) for _, _, idxs in self.buckets]
        self.launched = [False] * len(self.buckets)
        self.cuda = arena.device.type == "cuda"
        self.comm_stream = torch.cuda.Stream(arena.device) if self.cuda else None
        self.works = []
        self.active = False
        self.skip_comm = False      # measurement only: run the step without the all-reduces
        native.grad_ready_hooks.append(self.on_grad_ready)
        for p in arena.params:
            p.register_post_accumulate_grad_hook(self.on_grad_ready)

    def close(self):
        if self.on_grad_ready in native.grad_ready_hooks:
            native.grad_ready_hooks.remove(self.on_grad_ready)

    # ------------------------------------------------------------------------------
    def begin(self):
        self.remaining = [len(idxs) for _, _, idxs in self.buckets]
        self.launched = [False] * len(self.buckets)
        self.events = [None] * len(self.buckets)
        self.sub_events = {}
        self.order = []
        self.works = []
        self.active = True

    def on_grad_ready(self, p):
        if not self.active:
            return
        b = self.bucket_of.get(id(p))
        if b is None:
            return
        self.remaining[b] -= 1
        if self.remaining[b] == 0 and not self.launched[b]:
            self._launch(b)

    def _launch(self, b: int, tail: bool = False):
        """Reduce bucket b.  ``tail``: the bucket is reduced after backward has ended (nothing left to overlap
        with, nothing to disturb), so it goes out in sub-ranges of at most the bucket cap - each with its own
        event, so AdamW of sub-range i runs under the all-reduce of sub-range i+1 - on a wider grid."""
        self.launched[b] = True
        self.order.append(b)
        lo, hi, _ = self.buckets[b]
        if self.world == 1 or self.skip_comm:
            return
        if self.transport == "symm":
            ev = torch.cuda.Event()
            ev.record()
            self.comm_stream.wait_event(ev)
            spans = [(lo, hi)]
            if tail and (hi - lo) > self.cap:
                step = (self.cap + 7) // 8 * 8
                spans = [(a, min(a + step, hi)) for a in range(lo, hi, step)]
            # tail: compute has ended, so a wider grid costs nothing (channels [0, 128): the rs_adamw range is idle in DDP)
            max_blocks = self.tail_blocks if tail else self.nblocks
            with torch.cuda.stream(self.comm_stream):
                for a, z in spans:
                    nblocks = max(4, min(max_blocks, 128, ((z - a) * 4) >> 18))      # ~256 KiB of bucket per block
                    if self.stage_buf is not None and a % 8 == 0 and (z - a) % 8 == 0:
                        _lib.call("b200_allreduce_bf16", self.arena.grad.data_ptr(), self.stage_buf.ptrs_dev,
                                  self.stage_buf.mc_ptr, self.symm.flags.ptrs_dev, self.rank, self.world, a, z - a,
                                  1.0 / self.world, nblocks, CH_ALLREDUCE, self.comm_stream.cuda_stream)
                    else:
                        _lib.call("b200_allreduce", self.grad_buf.ptrs_dev, self.grad_buf.mc_ptr or None,
                                  self.symm.flags.ptrs_dev, self.rank, self.world, a, z - a, 1.0 / self.world,
                                  nblocks, CH_ALLREDUCE, self.comm_stream.cuda_stream)
                    e = torch.cuda.Event()
                    e.record(self.comm_stream)
                    self.sub_events.setdefault(b, []).append((a, z, e))
                self.events[b] = self.sub_events[b][-1][2]
        elif self.cuda:
            ev = torch.cuda.Event()
            ev.record()
            self.comm_stream.wait_event(ev)
            with torch.cuda.stream(self.comm_stream):
                view = self.arena.grad[lo:hi]
                view.mul_(1.0 / self.world)
                dist.all_reduce(view, group=self.group)
                self.events[b] = torch.cuda.Event()
                self.events[b].record(self.comm_stream)
        else:
            view = self.arena.grad[lo:hi]
            view.mul_(1.0 / self.world)
            self.works.append(dist.all_reduce(view, group=self.group, async_op=True))

    def finish(self, optimizer=None):
        """Reduce whatever was not triggered by a hook, join the communication stream and (when an
        optimizer with ranged steps is given) run AdamW bucket range by bucket range, so the update
        of everything already reduced overlaps the all-reduce of the last bucket (the input
        embedding, whose gradient only completes at the very end of backward)."""
        for b in range(len(self.buckets)):
            if not self.launched[b]:
                self._launch(b, tail=True)
        for w in self.works:
            w.wait()
        self.works = []
        self.active = False
        ranged = (optimizer is not None and getattr(optimizer, "native", False) and self.cuda and self.world > 1
                  and not self.skip_comm and len(self.order) > 1 and all(self.events[b] is not None for b in self.order))
        if not ranged:
            if self.cuda and self.world > 1 and not self.skip_comm:
                torch.cuda.current_stream().wait_stream(self.comm_stream)
            if optimizer is not None:
                optimizer.step()
            return
        cur = torch.cuda.current_stream()
        head, tail = self.order[:-1], self.order[-1]
        for b in head:
            cur.wait_event(self.events[b])
        # the head buckets are contiguous slices; step them as maximal contiguous ranges
        spans = sorted((self.buckets[b][0], self.buckets[b][1]) for b in head)
        merged = []
        for lo, hi in spans:
            if merged and merged[-1][1] == lo:
                merged[-1][1] = hi
            else:
                merged.append([lo, hi])
        first = True
        for lo, hi in merged:
            optimizer.step(lo, hi, tick=first)
            first = False
        for a, z, e in self.sub_events.get(tail) or [(self.buckets[tail][0], self.buckets[tail][1], self.events[tail])]:
            cur.wait_event(e)
            optimizer.step(a, z, tick=False)


class DDPStrategy:
    """Trainer strategy for main-ddp.py (and the dp dimension of main-pipe-ddp.py)."""

    def __init__(self, group=None):
        self.group = group
        self.comm_kind = None
        self._graphed = None

    def setup(self, model, args, device):
        self.device = torch.device(device)
        self.world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(self.group) if dist.is_initialized() else 0
        self.model = model.to(self.device)
        cuda = self.device.type == "cuda"
        want = getattr(args, "comm", "auto")
        native_ok = cuda and _lib.available() and not args.disable_amp
        transport = "gloo" if not cuda else "nccl"
        self.symm = None
        grad_buf = None
        if self.world > 1 and native_ok and want in ("auto", "symm") and symmetric_memory_available(self.device):
            try:
                self.symm = SymmContext(self.device, self.group)
                transport = "symm"
            except Exception as e:  # fabric / driver without peer mapping
                if want == "symm":
                    raise
                print(f"[ddp] symmetric memory unavailable ({e!r}); falling back to NCCL", flush=True)
        params = [p for p in self.model.parameters() if p.requires_grad]
        grad_alloc = None
        if transport == "symm":
            holder = {}

            def grad_alloc(n):
                holder["buf"] = self.symm.alloc(n, torch.float32)
                return holder["buf"].tensor
        self.arena = ParamArena(params, self.device, with_bf16=native_ok, with_moments=native_ok,
                                grad_alloc=grad_alloc)
        if transport == "symm":
            grad_buf = holder["buf"]
        self.optim = FusedAdamW(params, lr=args.learning_rate, device=self.device, arena=self.arena)
        self.comm_kind = transport
        self._verify_replicas()
        self.reducer = GradReducer(self.arena, self.group, getattr(args, "bucket_mb", 16.0), transport,
                                   self.symm, grad_buf, nblocks=int(os.environ.get("B200GPT_DDP_BLOCKS", "16")),
                                   comm_dtype=getattr(args, "comm_dtype", "fp32"))
        if transport == "symm":
            self.comm_kind = f"symm:{self.reducer.comm_dtype}"
        self.use_graph = (transport == "symm" and not getattr(args, "disable_cuda_graph", False)) or \
                         (self.world == 1 and native_ok and not getattr(args, "disable_cuda_graph", False))

    def _verify_replicas(self):
        """Replicas start identical because every rank seeds identically (no 122 MiB broadcast,
        SURVEY C2); one checksum all-reduce proves it, else rank 0's weights are broadcast."""
        if self.world == 1:
            return
        s = self.arena.fp32.double().sum().reshape(1)
        lo, hi = s.clone(), s.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=self.group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=self.group)
        if float(lo) != float(hi):
            src = dist.get_global_rank(self.group, 0) if self.group is not None else 0
            dist.broadcast(self.arena.fp32, src=src, group=self.group)
            self.arena.refresh_bf16()

    def train(self):
        self.model.train()

    def eval(self):
        self.model.eval()

    def _autocast(self, amp):
        if amp and self.device.type == "cuda":
            return torch.autocast(device_type="cuda", dtype=torch.bfloat16)
        return nullcontext()

    def _raw_step(self, input_ids, position_ids, targets, mask, amp):
        from ..utils import nvtx
        self.optim.zero_grad()
        self.reducer.begin()
        with self._autocast(amp), nvtx.phase("forward"):
            loss, _, _ = self.model.loss(input_ids, position_ids, targets, mask=mask)
        with nvtx.phase("backward + bucket all-reduce"):
            loss.backward()
        with nvtx.phase("grad-sync tail + optimizer"):
            self.reducer.finish(self.optim)       # joins communication and runs the (ranged) optimizer step
        return loss.detach()

    def optimizer_state(self):
        return self.optim.state_dict()

    def load_optimizer_state(self, sd):
        self.optim.load_state_dict(sd)

    def train_step(self, batch, targets, amp=True):
        if self.use_graph and amp:
            from ..runtime.graph import GraphedStep
            if self._graphed is None:
                self._graphed = GraphedStep(lambda i, p, t: self._raw_step(i, p, t, None, True))
            return self._graphed(batch["input_ids"], batch["position_ids"], targets)
        return self._raw_step(batch["input_ids"], batch["position_ids"], targets, batch.get("mask"), amp)

    @torch.no_grad()
    def eval_step(self, batch, targets, amp=True):
        with self._autocast(amp):
            return self.model.loss(batch["input_ids"], batch["position_ids"], targets, mask=batch.get("mask"))

    def all_reduce_mean(self, t):
        """One fused 2-float reduce per validation batch (reference: two, main-ddp.py:159-160)."""
        if self.world > 1:
            dist.all_reduce(t, group=self.group)
            t /= self.world
        return t

    def launches_per_step(self):
        return self._graphed.native_calls_per_replay if (self._graphed and self._graphed.graph) else None

    def generation_model(self):
        return self.model

    def close(self):
        self.reducer.close()
        self._graphed = None

    def state_dict(self):
        from ..utils import canonical_state_dict
        return canonical_state_dict(self.model)
