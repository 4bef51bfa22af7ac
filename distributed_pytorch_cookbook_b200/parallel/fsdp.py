"""Fully-sharded data parallel (ZeRO-3) engine.

Reference: ``FSDP(model, auto_wrap_policy=size_based(min_num_params=100), cpu_offload=...)``
(main-fsdp.py:60-69): every Linear / Embedding / LayerNorm becomes its own FSDP unit (68 + root at
the default config), each forward and backward all-gathers the unit's fp32 flat parameter and
each backward reduce-scatters its fp32 gradient - 196 x 3 NCCL collectives per step, most of
them tiny and latency bound.

This engine shards *per decoder layer* (plus an embeddings unit and a head unit), in two shard
groups per unit - ``big`` (matrices, consumed by the tensor-core kernels as bf16) and ``small``
(LayerNorm, biases, embeddings: consumed as fp32) - each split evenly over the ranks:

* every rank keeps fp32 master shards + AdamW moments (1/world of the model);
* forward needs a unit's full weights: the peers' **bf16** shards (fp32 for the small group) are
  pulled over NVLink straight into the local full buffer by ``b200_allgather`` on a side stream,
  one unit ahead of the compute stream (prefetch).  Gathered weights stay resident for the backward
  pass - 180 GB of HBM per GPU buys the second all-gather of the reference away;
* the kernels write weight gradients into a full-size fp32 gradient buffer that lives in symmetric
  memory; as soon as a unit's gradients are complete ``b200_rs_adamw`` runs on the side stream:
  the owner of each shard reduces it out of all peers' buffers (``multimem.ld_reduce`` when NVLS is
  available), applies 1/world, runs AdamW on its master shard and writes the bf16 shard the peers
  will gather next step.  Reduce-scatter, cast and optimizer are one kernel; no NCCL on the path.

Two memory modes.  *Resident* (default): the gathered bf16 weights and the full-size gradient buffer stay
allocated for the whole step, so backward needs no second all-gather - the fastest choice while the model fits
(10 B/param un-sharded + 18 B/param / world).  *Sharded* (``--reshard_after_forward``, the reference's
FULL_SHARD, main-fsdp.py:60-69): units are gathered into two rotating unit-sized buffers, freed for reuse as
soon as the next-but-one unit needs the slot, gathered again right before their backward (prefetched one unit
ahead), and their gradients live in a two-unit ring that is reduce-scattered into the shard-sized buffer - per
GPU that is 18 (+4) B/param / world plus two units, so the model size scales with the number of GPUs.

Fallback transports (``--comm nccl`` or CPU/gloo) run the same schedule with
``all_gather_into_tensor`` / ``all_reduce`` + the same AdamW maths in PyTorch.
``--cpu_offload`` keeps master shards and moments in pinned host memory and steps them on the CPU.
"""
from __future__ import annotations

import math
from contextlib import nullcontext
from typing import Dict, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from ..ops import _lib, native
from ..runtime.symm import (CH_ALLGATHER, CH_RS_ADAMW, SymmContext, symmetric_memory_available)

ALIGN = 128


def _round_up(n, m):
    return (n + m - 1) // m * m


def is_small(p: nn.Parameter, module_of: Dict[int, nn.Module]) -> bool:
    """fp32-consumed parameters: vectors (LayerNorm, biases) and embedding tables."""
    return p.ndim == 1 or isinstance(module_of.get(id(p)), nn.Embedding)


class ShardGroup:
    """A set of parameters flattened together and split evenly over the ranks."""

    def __init__(self, name: str, params: List[nn.Parameter], kind: str, world: int):
        self.name, self.params, self.kind, self.world = name, params, kind, world
        self.offsets, off = [], 0
        for p in params:
            self.offsets.append(off)
            off += _round_up(p.numel(), ALIGN)
        self.n = _round_up(max(off, ALIGN), ALIGN * world)
        self.shard_n = self.n // world
        self.base = 0            # offset of this group in the global flat space (set by the planner)
        self.ready = 0           # parameters whose gradient is complete in the current step
        self.gathered = False

    @property
    def shard_base(self) -> int:
        return self.base // self.world


def plan_units(model) -> List[List[nn.Module]]:
    """Sharding units in execution order: embeddings | each decoder layer | norm_out + lm_head."""
    return [[model.embeddings]] + [[layer] for layer in model.decoder.layers] + [[model.norm_out, model.lm_head]]


def plan_groups(model, world: int) -> List[List[ShardGroup]]:
    module_of = {}
    for m in model.modules():
        for p in m.parameters(recurse=False):
            module_of[id(p)] = m
    units, base = [], 0
    for ui, mods in enumerate(plan_units(model)):
        ps = [p for m in mods for p in m.parameters() if p.requires_grad]
        small = [p for p in ps if is_small(p, module_of)]
        big = [p for p in ps if not is_small(p, module_of)]
        groups = []
        for kind, lst in (("small", small), ("big", big)):
            if lst:
                g = ShardGroup(f"unit{ui}.{kind}", lst, kind, world)
                g.base = base
                base += g.n
                groups.append(g)
        units.append(groups)
    return units


def adamw_torch_(p, g, m, v, step: float, lr, beta1, beta2, eps, wd):
    """torch.optim.AdamW maths on flat shards (fallback / cpu_offload path)."""
    p.mul_(1.0 - lr * wd)
    m.mul_(beta1).add_(g, alpha=1.0 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1.0 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)


class FSDPStrategy:
    """Trainer strategy for main-fsdp.py."""

    def __init__(self, group=None, cpu_offload: bool = False):
        self.group, self.cpu_offload = group, cpu_offload
        self.comm_kind = None
        self._graphed = None
        self.betas, self.eps, self.wd = (0.9, 0.999), 1e-8, 0.01

    # ------------------------------------------------------------------------------
    def setup(self, model, args, device):
        self.device = torch.device(device)
        self.world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(self.group) if dist.is_initialized() else 0
        self.cpu_offload = self.cpu_offload or getattr(args, "cpu_offload", False)
        self.lr = args.learning_rate
        self.model = model.to(self.device)          # reference materialises the full model first too
        cuda = self.device.type == "cuda"
        self.native = cuda and _lib.available() and not args.disable_amp
        want = getattr(args, "comm", "auto")
        transport = "nccl" if cuda else "gloo"
        self.symm = None
        if (self.world > 1 and self.native and not self.cpu_offload and want in ("auto", "symm")
                and symmetric_memory_available(self.device)):
            try:
                self.symm = SymmContext(self.device, self.group)
                transport = "symm"
            except Exception as e:
                if want == "symm":
                    raise
                print(f"[fsdp] symmetric memory unavailable ({e!r}); falling back to NCCL", flush=True)
        self.comm_kind = transport
        self.units = plan_groups(self.model, self.world)
        self.groups = [g for u in self.units for g in u]
        self.G = sum(g.n for g in self.groups)
        Gs = self.G // self.world
        dev = self.device
        f32, bf16 = torch.float32, torch.bfloat16

        def alloc(n, dtype, symmetric):
            if symmetric and transport == "symm":
                buf = self.symm.alloc(n, dtype)
                return buf.tensor, buf
            return torch.zeros(n, dtype=dtype, device=dev), None

        import os
        self.reshard = bool(getattr(args, "reshard_after_forward", False)) or os.environ.get("B200GPT_FSDP_RESHARD") == "1"
        assert not (self.reshard and self.cpu_offload), "--reshard_after_forward and --cpu_offload are not combined"
        # "full space": where gathered weights and un-reduced gradients live.  Resident mode: the whole model
        # (offset = g.base).  Sharded mode: two unit-sized slots, unit u uses slot u % 2.
        for ui, unit in enumerate(self.units):
            off = 0
            for g in unit:
                g.unit, g.local = ui, off
                off += g.n
        if self.reshard:
            self.Umax = _round_up(max(sum(g.n for g in u) for u in self.units), ALIGN * self.world)
            for g in self.groups:
                g.fbase = (g.unit % 2) * self.Umax + g.local
            self.F = 2 * self.Umax
        else:
            for g in self.groups:
                g.fbase = g.base
            self.F = self.G
        self.slot_unit = [None, None]                # sharded mode: which unit's weights each slot holds
        self.slot_events = [[], []]                  # ... and the reduce-scatters still reading its gradient slot
        init_fp32 = torch.zeros(self.G, dtype=f32, device=dev) if self.reshard else None
        self.full_fp32 = torch.zeros(self.F, dtype=f32, device=dev)
        self.full_bf16 = torch.zeros(self.F, dtype=bf16, device=dev) if self.native else None
        self.grads, self.grads_buf = alloc(self.F, f32, True)
        state_dev = torch.device("cpu") if self.cpu_offload else dev
        if self.cpu_offload:
            pin = cuda
            self.master = torch.zeros(Gs, dtype=f32, pin_memory=pin)
            self.master_buf = None
            self.shard_grad_host = torch.zeros(Gs, dtype=f32, pin_memory=pin)
            self.master_dev = torch.zeros(Gs, dtype=f32, device=dev)     # staging for the gathers
        else:
            self.master, self.master_buf = alloc(Gs, f32, True)
            self.master_dev = self.master
        self.exp_avg = torch.zeros(Gs, dtype=f32, device=state_dev)
        self.exp_avg_sq = torch.zeros(Gs, dtype=f32, device=state_dev)
        self.shard_bf16, self.shard_bf16_buf = alloc(Gs, bf16, True) if self.native else (None, None)
        self.hyper = torch.tensor([self.lr, 0.0, 0.0, 0.0], dtype=f32, device=dev)
        self.host_step = 0
        # B200GPT_FSDP_RS=fused: the one-kernel reduce-scatter + AdamW (b200_rs_adamw, up to 64 resident blocks);
        # default: small-grid b200_reduce_scatter into a shard-sized buffer + b200_adamw behind it
        self.rs_split = transport == "symm" and os.environ.get("B200GPT_FSDP_RS", "split") != "fused"
        self.rs_blocks = int(os.environ.get("B200GPT_FSDP_RS_BLOCKS", "8"))
        self.shard_grad = torch.zeros(Gs, dtype=f32, device=dev) if (self.rs_split or transport == "nccl") else None

        # re-point parameters / gradients at the flat buffers and cut the shards
        self.group_of = {}
        self.param_names = {id(p): n for n, p in self.model.named_parameters()}
        with torch.no_grad():
            for g in self.groups:
                src = init_fp32 if self.reshard else self.full_fp32      # where the initial values are collected
                for p, o in zip(g.params, g.offsets):
                    n = p.numel()
                    lo = g.fbase + o
                    src[g.base + o:g.base + o + n].copy_(p.data.reshape(-1))
                    if self.reshard and self.native and g.kind == "big":
                        # consumed as bf16 only: the fp32 tensor keeps its shape but owns one element
                        p.data = torch.zeros(1, dtype=f32, device=dev).expand(p.shape)
                    else:
                        p.data = self.full_fp32[lo:lo + n].view(p.shape)
                    p.grad = self.grads[lo:lo + n].view(p.shape)
                    if self.native and g.kind == "big":
                        p._b200_bf16 = self.full_bf16[lo:lo + n].view(p.shape)
                        p._b200_bf16_version = p._version
                    self.group_of[id(p)] = g
                    p.register_post_accumulate_grad_hook(self._on_grad_ready)
                s_lo = g.base + self.rank * g.shard_n
                self.master[g.shard_base:g.shard_base + g.shard_n].copy_(src[s_lo:s_lo + g.shard_n])
                if self.native:
                    self.shard_bf16[g.shard_base:g.shard_base + g.shard_n].copy_(src[s_lo:s_lo + g.shard_n])
            del init_fp32
            if self.cpu_offload:
                self.master_dev.copy_(self.master, non_blocking=True)
        # fused all-gather: big (bf16) groups are not gathered by a separate kernel - the first GEMM that
        # consumes each weight pulls it from the peers' shards itself (b200_ag_gemm_bf16)
        # All-gather inside the consuming GEMM (b200_ag_gemm_bf16) or as a pull kernel on the side stream.  The
        # fused form runs on single-CTA tiles (it needs the gather warp), which costs more than it hides when the
        # weights are small: measured at 4 GPUs, default GPT 5.26 ms fused vs 5.09 ms unfused, 1.4 B model 90.6 vs
        # 91.5 ms.  B200GPT_FUSED_AG=1 / 0 forces it; unset, it is used for model widths >= 1024.
        fa = os.environ.get("B200GPT_FUSED_AG", "")
        wide = int(getattr(model, "dim", 0) or 0) >= 1024
        # (sharded mode re-gathers every unit for its backward on the side stream anyway: no fused gather there)
        self.fused_ag = transport == "symm" and (fa == "1" or (fa != "0" and wide)) and not self.reshard
        for g in self.groups:
            g.fused = False
        if self.fused_ag:
            head = self.units[-1]
            for g in self.groups:
                # the LM-head weight is consumed by the softmax-epilogue GEMM, which has no gather warp: its
                # group is pulled on the side stream one unit ahead like the small groups
                if g.kind != "big" or any(g is h for h in head):
                    continue
                g.fused = True
                g.peers_dev = torch.tensor([pp + 2 * g.shard_base for pp in self.shard_bf16_buf.peer_ptrs],
                                           dtype=torch.int64, device=dev)
                g.ag_sync = torch.zeros(2 + 1024, dtype=torch.int32, device=dev)
                for p, o in zip(g.params, g.offsets):
                    p._b200_ag = (self, g, 2 * o, 2 * p.numel())
                    p._b200_ag_version = 0
        native.grad_ready_hooks.append(self._on_grad_ready)
        # a unit's gradient buffer is only valid during that unit's backward: LayerNorm hand-offs stay inside a block
        native.set_fuse_scope("block")
        self.comm_stream = torch.cuda.Stream(dev) if cuda else None
        self.version = 1
        for g in self.groups:
            g.gathered_version = 0
            g.event = None
        self.active = False
        self.skip_comm = False
        # forward pre-hooks: make the unit's weights resident (and prefetch the next unit)
        unit_mods = plan_units(self.model)
        for ui, mods in enumerate(unit_mods):
            mods[0].register_forward_pre_hook(lambda m, inp, ui=ui: self._pre_forward(ui))
            if self.reshard and ui < len(unit_mods) - 1:
                # sharded mode: the gradient of a unit's output arriving is the cue to bring its weights back
                mods[-1].register_forward_hook(lambda m, inp, out, ui=ui: self._arm_backward(ui, out))
        # the native path reads norm_out / lm_head weights without calling those modules (models/gpt.py head_input,
        # loss), so their forward pre-hook never fires there: hook the head unit on head_input itself
        head_ui = len(unit_mods) - 1
        orig_head_input = self.model.head_input

        def head_input(x, _orig=orig_head_input, _ui=head_ui):
            self._pre_forward(_ui)
            return _orig(x)
        self.model.head_input = head_input
        if cuda:
            torch.cuda.synchronize(dev)
        if dist.is_initialized():
            dist.barrier(group=self.group)
        self.use_graph = transport == "symm" and not getattr(args, "disable_cuda_graph", False)

    # ------------------------------------------------------------------------------
    def _gather_dtype(self, g: ShardGroup):
        return torch.bfloat16 if (self.native and g.kind == "big") else torch.float32

    def _gather(self, g: ShardGroup, dtype=None, dest=None):
        """All-gather one shard group into the local full buffer (on the side stream on CUDA), or into ``dest``."""
        dtype = dtype or self._gather_dtype(g)
        esz = 2 if dtype == torch.bfloat16 else 4
        full = self.full_bf16 if dtype == torch.bfloat16 else self.full_fp32
        shard = self.shard_bf16 if dtype == torch.bfloat16 else self.master_dev
        full_view = dest if dest is not None else full[g.fbase:g.fbase + g.n]
        shard_view = shard[g.shard_base:g.shard_base + g.shard_n]
        if self.world == 1:
            full_view.copy_(shard_view)
            return
        if self.skip_comm:                                # measurement only: no traffic, the weights go stale
            return
        cuda = self.device.type == "cuda"
        ctx = nullcontext()
        if cuda:
            ev = torch.cuda.Event()
            ev.record()                                   # compute is done with the old weights
            self.comm_stream.wait_event(ev)
            ctx = torch.cuda.stream(self.comm_stream)
        with ctx:
            if self.comm_kind == "symm":
                buf = self.shard_bf16_buf if dtype == torch.bfloat16 else self.master_buf
                nblocks = max(2, min(64, (g.shard_n * esz * self.world) >> 18))   # ~256 KiB in flight per block
                _lib.call("b200_allgather", buf.ptrs_dev, self.symm.flags.ptrs_dev, self.rank, self.world,
                          g.shard_base * esz, g.shard_n * esz, full_view.data_ptr(), 1, nblocks, CH_ALLGATHER,
                          self.comm_stream.cuda_stream)
            elif cuda:
                dist.all_gather_into_tensor(full_view, shard_view.contiguous(), group=self.group)
            else:
                parts = list(full_view.view(self.world, g.shard_n).unbind(0))
                dist.all_gather(parts, shard_view.contiguous(), group=self.group)
            if cuda:
                g.event = torch.cuda.Event()
                g.event.record(self.comm_stream)

    def _ensure(self, g: ShardGroup):
        if g.fused:
            return                                        # gathered inside the consuming GEMMs
        if self.reshard:
            slot = g.unit % 2
            if self.slot_unit[slot] != g.unit:            # the slot held the unit two places away: evict it
                if self.slot_unit[slot] is not None:
                    for h in self.units[self.slot_unit[slot]]:
                        h.gathered_version = 0
                self.slot_unit[slot] = g.unit
        if g.gathered_version != self.version:
            self._gather(g)
            g.gathered_version = self.version

    def _wait_unit(self, ui: int):
        if self.device.type == "cuda":
            for g in self.units[ui]:
                if g.event is not None:
                    torch.cuda.current_stream().wait_event(g.event)

    def _pre_forward(self, ui: int):
        for g in self.units[ui]:
            self._ensure(g)
        if ui + 1 < len(self.units):                      # prefetch the next unit under this one's compute
            for g in self.units[ui + 1]:
                self._ensure(g)
        self._wait_unit(ui)

    # -- sharded mode: weights come back, and the gradient slot is recycled, right before a unit's backward --------
    def _arm_backward(self, ui: int, out):
        t = out[0] if isinstance(out, (tuple, list)) else out
        if self.active and torch.is_tensor(t) and t.requires_grad:
            t.register_hook(lambda grad, ui=ui: self._pre_backward(ui))

    def _pre_backward(self, ui: int):
        if not self.active or self.bwd_seen.get(ui):
            return None
        self.bwd_seen[ui] = True
        cuda = self.device.type == "cuda"
        # this unit's gradient slot last held unit ui + 2: its reduce-scatters must be done reading, then clear it
        slot = ui % 2
        if ui + 2 < len(self.units):
            if cuda:
                for e in self.slot_events[slot]:
                    torch.cuda.current_stream().wait_event(e)
            self.slot_events[slot] = []
            self.grads[slot * self.Umax:(slot + 1) * self.Umax].zero_()
        for g in self.units[ui]:
            self._ensure(g)
        if ui > 0:                                        # prefetch the unit whose backward comes next
            for g in self.units[ui - 1]:
                self._ensure(g)
        self._wait_unit(ui)
        return None

    # ------------------------------------------------------------------------------
    def _on_grad_ready(self, p):
        if not self.active:
            return
        g = self.group_of.get(id(p))
        if g is None:
            return
        g.ready += 1
        if g.ready == len(g.params):
            self._reduce_and_step(g)

    def _reduce_and_step(self, g: ShardGroup):
        g.ready = -(1 << 30)                               # launched
        b1, b2 = self.betas
        cuda = self.device.type == "cuda"
        sb, sn = g.shard_base, g.shard_n
        lo = g.fbase + self.rank * sn
        ctx = nullcontext()
        if cuda:
            ev = torch.cuda.Event()
            ev.record()
            self.comm_stream.wait_event(ev)
            ctx = torch.cuda.stream(self.comm_stream)
        with ctx:
            self._reduce_and_step_on_stream(g, sb, sn, lo, b1, b2, cuda)
            if cuda and self.reshard:
                e = torch.cuda.Event()
                e.record(self.comm_stream)
                self.slot_events[g.unit % 2].append(e)

    def _reduce_and_step_on_stream(self, g, sb, sn, lo, b1, b2, cuda):
        if True:
            if self.skip_comm:                            # measurement only: AdamW on the shard's local gradients
                pb = self.shard_bf16.data_ptr() + 2 * sb if g.kind == "big" else None
                _lib.call("b200_adamw", self.master.data_ptr() + 4 * sb, self.grads.data_ptr() + 4 * lo,
                          self.exp_avg.data_ptr() + 4 * sb, self.exp_avg_sq.data_ptr() + 4 * sb, pb, sn,
                          self.hyper.data_ptr(), b1, b2, self.eps, self.wd, self.comm_stream.cuda_stream)
                return
            if self.comm_kind == "symm" and self.rs_split:
                # reduce-scatter on a small grid (the switch adds; <= 16 blocks leave the SMs to the compute
                # kernels), then AdamW on the reduced shard as an ordinary kernel behind it
                pb = self.shard_bf16.data_ptr() + 2 * sb if g.kind == "big" else None
                nblocks = max(1, min(self.rs_blocks, (sn * 4) >> 20))           # ~1 MiB of shard per block
                st = self.comm_stream.cuda_stream
                _lib.call("b200_reduce_scatter", self.grads_buf.ptrs_dev, self.grads_buf.mc_ptr or None,
                          self.symm.flags.ptrs_dev, self.rank, self.world, lo, sn, 1.0 / self.world,
                          self.shard_grad.data_ptr() + 4 * sb, 1, nblocks, CH_RS_ADAMW, st)
                _lib.call("b200_adamw", self.master.data_ptr() + 4 * sb, self.shard_grad.data_ptr() + 4 * sb,
                          self.exp_avg.data_ptr() + 4 * sb, self.exp_avg_sq.data_ptr() + 4 * sb, pb, sn,
                          self.hyper.data_ptr(), b1, b2, self.eps, self.wd, st)
                return
            if self.comm_kind == "symm":
                pb = self.shard_bf16.data_ptr() + 2 * sb if g.kind == "big" else None
                nblocks = max(2, min(64, (sn * 4) >> 17))                        # ~128 KiB of shard per block
                _lib.call("b200_rs_adamw", self.grads_buf.ptrs_dev, self.grads_buf.mc_ptr or None,
                          self.symm.flags.ptrs_dev, self.rank, self.world, lo, sn, 1.0 / self.world,
                          self.master.data_ptr() + 4 * sb, self.exp_avg.data_ptr() + 4 * sb,
                          self.exp_avg_sq.data_ptr() + 4 * sb, pb, self.hyper.data_ptr(), b1, b2, self.eps,
                          self.wd, 1, nblocks, CH_RS_ADAMW, self.comm_stream.cuda_stream)
                return
            gview = self.grads[g.fbase:g.fbase + g.n]
            gs = self.grads[lo:lo + sn]
            if self.world > 1:
                gview.mul_(1.0 / self.world)
                if self.comm_kind == "nccl" and self.shard_grad is not None:
                    gs = self.shard_grad[sb:sb + sn]       # the reference's collective: reduce-scatter into the shard
                    dist.reduce_scatter_tensor(gs, gview, group=self.group)
                else:
                    dist.all_reduce(gview, group=self.group)   # gloo has no reduce_scatter
            if self.cpu_offload:
                self.shard_grad_host[sb:sb + sn].copy_(gs, non_blocking=False)
                adamw_torch_(self.master[sb:sb + sn], self.shard_grad_host[sb:sb + sn], self.exp_avg[sb:sb + sn],
                             self.exp_avg_sq[sb:sb + sn], float(self.host_step), self.lr, b1, b2, self.eps, self.wd)
                self.master_dev[sb:sb + sn].copy_(self.master[sb:sb + sn], non_blocking=True)
            else:
                adamw_torch_(self.master[sb:sb + sn], gs, self.exp_avg[sb:sb + sn], self.exp_avg_sq[sb:sb + sn],
                             float(self.host_step), self.lr, b1, b2, self.eps, self.wd)
            if self.native:
                self.shard_bf16[sb:sb + sn].copy_(self.master_dev[sb:sb + sn])

    # ------------------------------------------------------------------------------
    def train(self):
        self.model.train()

    def eval(self):
        self.model.eval()

    def _autocast(self, amp):
        if amp and self.device.type == "cuda":
            return torch.autocast(device_type="cuda", dtype=torch.bfloat16)
        return nullcontext()

    def _raw_step(self, input_ids, position_ids, targets, mask, amp):
        self.version += 1                                  # weights of the previous step are stale
        self.host_step += 1
        self.grads.zero_()
        if self.native:
            _lib.call("b200_adamw_tick", self.hyper.data_ptr(), _lib.stream())
        else:
            self.hyper[1] += 1
        for g in self.groups:
            g.ready = 0
        self.bwd_seen = {}
        self.slot_events = [[], []]
        self.active = True
        from ..utils import nvtx
        with self._autocast(amp), nvtx.phase("forward (+ all-gather)"):
            loss, _, _ = self.model.loss(input_ids, position_ids, targets, mask=mask)
        with nvtx.phase("backward (+ reduce-scatter + AdamW)"):
            loss.backward()
        for g in self.groups:                              # parameters that received no gradient
            if g.ready >= 0:
                self._reduce_and_step(g)
        self.active = False
        if self.device.type == "cuda":
            if self.rs_split and self.fused_ag and not self.skip_comm:
                # the next step's GEMMs pull the peers' bf16 shards without a barrier of their own: every rank's
                # AdamW kernels (queued on the side stream above) must have finished first
                with torch.cuda.stream(self.comm_stream):
                    self.symm.barrier()
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        self.version += 1                                  # shards changed: gathered copies are stale
        return loss.detach()

    def train_step(self, batch, targets, amp=True):
        if self.use_graph and amp:
            from ..runtime.graph import GraphedStep
            if self._graphed is None:
                self._graphed = GraphedStep(lambda i, p, t: self._raw_step(i, p, t, None, True))
            out = self._graphed(batch["input_ids"], batch["position_ids"], targets)
            self.version += 2
            return out
        return self._raw_step(batch["input_ids"], batch["position_ids"], targets, batch.get("mask"), amp)

    @torch.no_grad()
    def eval_step(self, batch, targets, amp=True):
        with self._autocast(amp):
            return self.model.loss(batch["input_ids"], batch["position_ids"], targets, mask=batch.get("mask"))

    def all_reduce_mean(self, t):
        if self.world > 1:
            dist.all_reduce(t, group=self.group)
            t /= self.world
        return t

    def launches_per_step(self):
        return self._graphed.native_calls_per_replay if (self._graphed and self._graphed.graph) else None

    def set_skip_comm(self, on: bool):
        """Measurement aid (bench: exposed communication): run the same step with every collective left out -
        no gathers (weights go stale), AdamW on the local gradients of the own shard, plain CTA-pair GEMMs.
        Symmetric-memory transport on CUDA only; the captured graph is dropped and re-captured."""
        assert self.comm_kind == "symm", "skip_comm is a measurement aid of the symmetric-memory transport"
        self.skip_comm = on
        for g in self.groups:
            g.event = None                                # events recorded inside the dropped graph capture are dead
            if hasattr(g, "peers_dev"):
                g.fused = self.fused_ag and not on
        self._graphed = None

    def optimizer_state(self):
        """This rank's shard of the optimizer state (fp32 master shard, moments, step)."""
        return {"fsdp": True, "world": self.world, "rank": self.rank, "hyper": self.hyper.cpu(),
                "host_step": self.host_step, "master": self.master.detach().cpu().clone(),
                "exp_avg": self.exp_avg.detach().cpu().clone(), "exp_avg_sq": self.exp_avg_sq.detach().cpu().clone()}

    def load_optimizer_state(self, sd):
        if not sd.get("fsdp") or sd["world"] != self.world or sd["rank"] != self.rank:
            raise ValueError("FSDP optimizer state belongs to a different rank / world size")
        self.hyper.copy_(sd["hyper"])
        self.hyper[0] = self.lr
        self.host_step = int(sd["host_step"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])

    def close(self):
        """Detach from the process-wide gradient-ready hook list (benchmarks run several strategies in one process)."""
        if self._on_grad_ready in native.grad_ready_hooks:
            native.grad_ready_hooks.remove(self._on_grad_ready)
            native.set_fuse_scope("all")
        self._graphed = None

    # ------------------------------------------------------------------------------
    def gather_full_fp32(self):
        """Collective: bring every fp32 master shard into ``param.data`` (full model on every rank) -
        the reference's FULL_STATE_DICT gather (main-fsdp.py:194)."""
        if self.device.type == "cuda":
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        for g in self.groups:
            self._gather(g, torch.float32)
            if self._gather_dtype(g) == torch.float32:
                g.gathered_version = self.version
            elif not g.fused:
                # the bf16 copy too: a forward that only some ranks run afterwards (rank-0 generation) must not
                # find a stale group and start a collective gather on its own
                self._gather(g)
                g.gathered_version = self.version
        if self.device.type == "cuda":
            torch.cuda.current_stream().wait_stream(self.comm_stream)

    def _full_state(self):
        """Sharded mode, collective: the canonical fp32 state_dict assembled from a temporary full-size buffer
        (the only moment a whole model's worth of fp32 exists on a GPU in this mode)."""
        from collections import OrderedDict
        if self.device.type == "cuda":
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        flat = torch.empty(self.G, dtype=torch.float32, device=self.device)
        for g in self.groups:
            self._gather(g, torch.float32, dest=flat[g.base:g.base + g.n])
        if self.device.type == "cuda":
            torch.cuda.current_stream().wait_stream(self.comm_stream)
        by_name = {}
        for g in self.groups:
            for p, o in zip(g.params, g.offsets):
                by_name[self.param_names[id(p)]] = flat[g.base + o:g.base + o + p.numel()].view(p.shape).cpu().clone()
        return OrderedDict((k, by_name[k]) for k, _ in self.model.named_parameters())

    def generation_model(self):
        """All ranks gather (collective) so that rank 0 can run the un-sharded model - avoids the
        reference's rank-0-only forward through FSDP, which deadlocks."""
        if self.reshard:
            sd = self._full_state()
            if self.rank != 0:
                return None
            if getattr(self, "_gen_model", None) is None:
                m = self.model
                self._gen_model = type(m)(dim=m.dim, head_dim=m.head_dim, heads=m.heads, num_layers=m.num_layers,
                                          vocab_size=m.vocab_size,
                                          max_position_embeddings=m.max_position_embeddings).to(self.device)
            self._gen_model.load_state_dict(sd)
            for p in self._gen_model.parameters():
                if hasattr(p, "_b200_bf16_version"):
                    p._b200_bf16_version = -1
            return self._gen_model
        self.gather_full_fp32()
        return self.model

    def state_dict(self):
        from ..utils import canonical_state_dict
        if self.reshard:
            return self._full_state()
        self.gather_full_fp32()
        return canonical_state_dict(self.model)

    def memory_bytes(self):
        """Bytes this rank holds for parameters / gradients / optimizer state (what --reshard_after_forward shrinks)."""
        def nbytes(*ts):
            return sum(t.numel() * t.element_size() for t in ts if t is not None)
        return {"gathered_params": nbytes(self.full_fp32, self.full_bf16), "gradients": nbytes(self.grads, self.shard_grad),
                "shards_and_optimizer": nbytes(self.master, self.shard_bf16, self.exp_avg, self.exp_avg_sq),
                "mode": "sharded" if self.reshard else "resident", "world": self.world, "params": self.G}
