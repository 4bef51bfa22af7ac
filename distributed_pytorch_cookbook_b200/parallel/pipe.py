"""Pipeline-parallel engine: layer partitioner, stage module, GPipe / 1F1B schedules over p2p.

Reference intent (main-pipe.py:52-83, which does not parse - SURVEY 2.8 R6): split the decoder
layers over ``num_stages`` devices, embeddings on the first stage, ``norm_out`` + ``lm_head`` on
the last, ``chunks = num_stages`` micro-batches, GPipe fill-drain inside one process with peer
copies.  Here: one process per stage (torchrun), activations / activation-gradients move with
batched ``isend``/``irecv`` (NCCL over NVLink on GPU boxes, gloo on CPU) in bf16 on the wire,
and both GPipe and 1F1B are the same loop with a different warm-up depth.  Each stage owns a
fused AdamW over its own parameters; the last stage evaluates the fused LM-head/CE on its own
copy of the targets (every rank iterates the same data, so nothing but activations is sent).
With ``dp_group`` set, stages are replicated data-parallel (main-pipe-ddp.py): the stage's
gradient arena is reduced over the dp group by the same bucketed reducer as DDP.
"""
from __future__ import annotations

from contextlib import nullcontext
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist
import torch.nn as nn

from ..ops import _lib
from ..optim import FusedAdamW
from ..runtime.arena import ParamArena


# ----------------------------------------------------------------------------------------
def partition_layers(num_layers: int, num_stages: int) -> List[int]:
    """Layers per stage, the reference's rule (main-pipe.py:57-65): with q, r = divmod(L, S) the
    first stage gets q + (r == S-1), middle stage i gets q + (i < r), the last stage q (it also
    holds norm_out + lm_head, so it is the lightest)."""
    L, S = num_layers, num_stages
    if S == 1:
        return [L]
    q, r = divmod(L, S)
    out = [q + int(r == S - 1)] + [q + int(i < r) for i in range(S - 2)] + [q]
    assert sum(out) == L, (out, L)
    return out


def stage_layer_range(num_layers: int, num_stages: int, stage: int) -> Tuple[int, int]:
    per = partition_layers(num_layers, num_stages)
    lo = sum(per[:stage])
    return lo, lo + per[stage]


class PipelineStage(nn.Module):
    """The slice of a ``TransformerDecoderLM`` owned by one stage.  Sub-modules keep their
    original names (``embeddings``, ``decoder.layers.<global index>``, ``norm_out``, ``lm_head``)
    so the union of all stages' state_dicts is the canonical checkpoint."""

    def __init__(self, full_model, stage: int, num_stages: int):
        super().__init__()
        self.stage, self.num_stages = stage, num_stages
        self.is_first, self.is_last = stage == 0, stage == num_stages - 1
        self.lo, self.hi = stage_layer_range(len(full_model.decoder.layers), num_stages, stage)
        self.dim = full_model.dim
        self.vocab_size = full_model.vocab_size
        self.embeddings = full_model.embeddings if self.is_first else None
        self.layers = nn.ModuleDict({str(i): full_model.decoder.layers[i] for i in range(self.lo, self.hi)})
        self.norm_out = full_model.norm_out if self.is_last else None
        self.lm_head = full_model.lm_head if self.is_last else None
        object.__setattr__(self, "_full", full_model)   # not a sub-module: avoids double registration

    def canonical_items(self):
        """(canonical key, tensor) pairs of the parameters this stage owns."""
        out = []
        if self.is_first:
            out += [("embeddings." + k, v) for k, v in self.embeddings.state_dict().items()]
        for i in range(self.lo, self.hi):
            out += [(f"decoder.layers.{i}." + k, v) for k, v in self.layers[str(i)].state_dict().items()]
        if self.is_last:
            out += [("norm_out." + k, v) for k, v in self.norm_out.state_dict().items()]
            out += [("lm_head." + k, v) for k, v in self.lm_head.state_dict().items()]
        return out

    def forward_hidden(self, x, input_ids=None, position_ids=None, mask=None):
        if self.is_first:
            x = self.embeddings(input_ids, position_ids)
        for i in range(self.lo, self.hi):
            x = self.layers[str(i)](x, mask)
        return x

    def loss(self, x, targets, grad_scale: float = 1.0):
        from .. import ops
        h = self._full.head_input(x)
        return ops.lmhead_cross_entropy(h, self.lm_head.weight, targets, grad_scale=grad_scale)

    def logits(self, x):
        from .. import ops
        h = self._full.head_input(x)
        if ops.use_native(h):
            return ops.linear(h, self.lm_head.weight, None)
        return self.lm_head(h)


# ----------------------------------------------------------------------------------------
def schedule_actions(kind: str, stage: int, num_stages: int, num_microbatches: int) -> List[Tuple[str, int]]:
    """Per-stage action list [('F', mb) | ('B', mb)] for GPipe or 1F1B (used for simulation in
    tests; the executor below follows the same order)."""
    M = num_microbatches
    warm = M if kind == "gpipe" else min(num_stages - stage - 1, M)
    acts = [("F", i) for i in range(warm)]
    f, b = warm, 0
    for _ in range(M - warm):
        acts.append(("F", f)); f += 1
        acts.append(("B", b)); b += 1
    acts += [("B", i) for i in range(b, M)]
    return acts


class _P2P:
    """Batched isend/irecv between neighbouring stages (ranks given as global ranks)."""

    def __init__(self, prev_rank: Optional[int], next_rank: Optional[int], shape, wire_dtype, device, group=None):
        self.prev, self.next, self.shape, self.dtype, self.device, self.group = \
            prev_rank, next_rank, tuple(shape), wire_dtype, device, group

    def _buf(self):
        return torch.empty(self.shape, dtype=self.dtype, device=self.device)

    def _run(self, ops):
        if not ops:
            return
        for w in dist.batch_isend_irecv(ops):
            w.wait()

    def exchange(self, send_next=None, send_prev=None, recv_prev=False, recv_next=False):
        ops, from_prev, from_next = [], None, None
        if send_prev is not None and self.prev is not None:
            ops.append(dist.P2POp(dist.isend, send_prev.to(self.dtype).contiguous(), self.prev, self.group))
        if recv_prev and self.prev is not None:
            from_prev = self._buf()
            ops.append(dist.P2POp(dist.irecv, from_prev, self.prev, self.group))
        if send_next is not None and self.next is not None:
            ops.append(dist.P2POp(dist.isend, send_next.to(self.dtype).contiguous(), self.next, self.group))
        if recv_next and self.next is not None:
            from_next = self._buf()
            ops.append(dist.P2POp(dist.irecv, from_next, self.next, self.group))
        self._run(ops)
        return from_prev, from_next


class PipeStrategy:
    """Trainer strategy for main-pipe.py (dp_size == 1) and main-pipe-ddp.py (dp_size > 1).

    Rank layout for pp x dp: ``rank = pp_rank * dp_size + dp_rank`` - neighbouring stages of one
    replica are ``dp_size`` ranks apart, the replicas of one stage are consecutive ranks.
    """

    def __init__(self, num_stages: int = 0, num_microbatches: int = 0, schedule: str = "gpipe", dp_size: int = 1):
        self.num_stages, self.num_microbatches, self.schedule, self.dp_size = \
            num_stages, num_microbatches, schedule, dp_size
        self.comm_kind = None

    # ------------------------------------------------------------------------------
    def setup(self, model, args, device):
        self.device = torch.device(device)
        world, rank = dist.get_world_size(), dist.get_rank()
        S = self.num_stages or (world // self.dp_size)
        assert S * self.dp_size == world, f"world {world} != stages {S} x dp {self.dp_size}"
        self.S, self.M = S, (self.num_microbatches or S)       # reference: chunks = num_stages
        self.pp_rank, self.dp_rank = rank // self.dp_size, rank % self.dp_size
        self.args = args
        # process groups: every rank must create every group, in the same order
        self.dp_group = None
        if self.dp_size > 1:
            for s in range(S):
                g = dist.new_group([s * self.dp_size + d for d in range(self.dp_size)])
                if s == self.pp_rank:
                    self.dp_group = g
        self.stage = PipelineStage(model, self.pp_rank, S).to(self.device)
        cuda = self.device.type == "cuda"
        self.native_ok = cuda and _lib.available() and not args.disable_amp
        params = [p for p in self.stage.parameters() if p.requires_grad]
        self.reducer = None
        symm = None
        grad_alloc, holder = None, {}
        transport = "gloo" if not cuda else "nccl"
        if self.dp_size > 1:
            from ..runtime.symm import SymmContext, symmetric_memory_available
            want = getattr(args, "comm", "auto")
            if self.native_ok and want in ("auto", "symm") and symmetric_memory_available(self.device):
                try:
                    symm = SymmContext(self.device, self.dp_group)
                    transport = "symm"

                    def grad_alloc(n):
                        holder["buf"] = symm.alloc(n, torch.float32)
                        return holder["buf"].tensor
                except Exception as e:
                    if want == "symm":
                        raise
                    print(f"[pipe-ddp] symmetric memory unavailable ({e!r}); using NCCL", flush=True)
        self.comm_kind = f"p2p:{'nccl' if cuda else 'gloo'}" + (f"+dp:{transport}" if self.dp_size > 1 else "")
        self.arena = ParamArena(params, self.device, with_bf16=self.native_ok, with_moments=self.native_ok,
                                grad_alloc=grad_alloc)
        self.optim = FusedAdamW(params, lr=args.learning_rate, device=self.device, arena=self.arena)
        if self.dp_size > 1:
            from .ddp import GradReducer
            self.reducer = GradReducer(self.arena, self.dp_group, getattr(args, "bucket_mb", 4.0), transport,
                                       symm, holder.get("buf"))
        self.mb = args.batch_size // self.M
        assert self.mb * self.M == args.batch_size, "batch_size must be divisible by the number of micro-batches"
        T = args.sequence_length - 1
        wire = torch.bfloat16 if (cuda and not args.disable_amp) else torch.float32
        prev_rank = rank - self.dp_size if self.pp_rank > 0 else None
        next_rank = rank + self.dp_size if self.pp_rank < S - 1 else None
        self.p2p = _P2P(prev_rank, next_rank, (self.mb, T, model.dim), wire, self.device)
        self.has_loss = self.stage.is_last

    def train(self):
        self.stage.train()

    def eval(self):
        self.stage.eval()

    def _autocast(self, amp):
        if amp and self.device.type == "cuda":
            return torch.autocast(device_type="cuda", dtype=torch.bfloat16)
        return nullcontext()

    # ------------------------------------------------------------------------------
    def _micro(self, t, i):
        return None if t is None else t[i * self.mb:(i + 1) * self.mb]

    def _forward(self, i, recv, batch, targets, amp, train=True):
        st = self.stage
        with self._autocast(amp):
            if st.is_first:
                x_in = None
                x = st.forward_hidden(None, self._micro(batch["input_ids"], i), self._micro(batch["position_ids"], i),
                                      self._micro(batch.get("mask"), i))
            else:
                x_in = recv.float()
                if train:
                    x_in.requires_grad_(True)
                x = st.forward_hidden(x_in, mask=self._micro(batch.get("mask"), i))
            if st.is_last:
                loss, n_valid, n_correct = st.loss(x, self._micro(targets, i), grad_scale=1.0 / self.M)
                return x_in, loss, (loss.detach(), n_valid, n_correct)
        return x_in, x, None

    def _backward(self, x_in, out, grad_from_next):
        if self.stage.is_last:
            (out * (1.0 / self.M)).backward()
        else:
            torch.autograd.backward(out, grad_from_next.to(out.dtype))
        return None if x_in is None else x_in.grad

    def train_step(self, batch, targets, amp=True):
        M, S = self.M, self.S
        self.optim.zero_grad()
        if self.reducer is not None:
            self.reducer.begin()
            self.reducer.active = False              # grads accumulate over micro-batches: reduce at the end
        warm = M if self.schedule == "gpipe" else min(S - self.pp_rank - 1, M)
        remaining = M - warm
        inflight, stats = [], []
        f = 0
        recv = None
        for _ in range(warm):
            recv, _ = self.p2p.exchange(recv_prev=True)
            x_in, out, st = self._forward(f, recv, batch, targets, amp)
            self.p2p.exchange(send_next=None if self.stage.is_last else out.detach())
            inflight.append((x_in, out)); stats.append(st); f += 1
        if remaining > 0:
            recv, _ = self.p2p.exchange(recv_prev=True)
        for k in range(remaining):
            x_in, out, st = self._forward(f, recv, batch, targets, amp)
            inflight.append((x_in, out)); stats.append(st); f += 1
            _, g = self.p2p.exchange(send_next=None if self.stage.is_last else out.detach(), recv_next=True)
            bx, bout = inflight.pop(0)
            gin = self._backward(bx, bout, g)
            if k == remaining - 1:
                self.p2p.exchange(send_prev=gin)
            else:
                recv, _ = self.p2p.exchange(send_prev=gin, recv_prev=True)
        while inflight:
            _, g = self.p2p.exchange(recv_next=True)
            bx, bout = inflight.pop(0)
            gin = self._backward(bx, bout, g)
            self.p2p.exchange(send_prev=gin)
        if self.reducer is not None:
            self.reducer.active = True
            self.reducer.finish()
        self.optim.step()
        if self.stage.is_last:
            loss = torch.stack([s[0].float() for s in stats]).mean()
        else:
            loss = torch.zeros((), device=self.device)
        # every rank reports the same loss (the trainer prints on rank 0): broadcast from the last stage
        src = (S - 1) * self.dp_size + self.dp_rank
        if self.dp_size == 1:
            dist.broadcast(loss, src=src)
        else:
            dist.all_reduce(loss)
            loss = loss / self.dp_size
        return loss.detach()

    @torch.no_grad()
    def eval_step(self, batch, targets, amp=True):
        stats = []
        for i in range(self.M):
            recv, _ = self.p2p.exchange(recv_prev=True)
            _, out, st = self._forward(i, recv, batch, targets, amp, train=False)
            self.p2p.exchange(send_next=None if self.stage.is_last else out)
            stats.append(st)
        res = torch.zeros(3, device=self.device)
        if self.stage.is_last:
            res[0] = torch.stack([s[0].float() for s in stats]).mean()
            res[1] = torch.stack([s[1].float() for s in stats]).sum()
            res[2] = torch.stack([s[2].float() for s in stats]).sum()
        dist.all_reduce(res)                      # only last-stage ranks contribute
        res = res / self.dp_size if self.dp_size > 1 else res
        return res[0], res[1], res[2]

    def all_reduce_mean(self, t):
        return t                                   # eval_step already returned group-wide values

    def close(self):
        if self.reducer is not None:
            self.reducer.close()

    static_batches = True      # micro-batches travel between stages in fixed-shape buffers

    def optimizer_state(self):
        return self.optim.state_dict()

    def load_optimizer_state(self, sd):
        self.optim.load_state_dict(sd)

    def generation_model(self):
        return None                                # generation needs all stages; see gather_full_model()

    # ------------------------------------------------------------------------------
    def state_dict(self):
        """Canonical checkpoint: every stage sends its (key, tensor) pairs to rank 0 (collective)."""
        mine = [(k, v.detach().float().cpu()) for k, v in self.stage.canonical_items()] if self.dp_rank == 0 else []
        gathered = [None] * dist.get_world_size() if dist.get_rank() == 0 else None
        dist.gather_object(mine, gathered, dst=0)
        if dist.get_rank() != 0:
            return None
        from collections import OrderedDict
        flat = dict(kv for part in gathered for kv in part)
        order = list(self.stage._full.state_dict().keys())
        return OrderedDict((k, flat[k]) for k in order)
