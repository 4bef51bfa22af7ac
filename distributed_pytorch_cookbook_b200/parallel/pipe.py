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


class _Link:
    """One direction of one stage boundary.

    Every boundary gets two 2-rank process groups, one per direction (activations forward, gradients
    backward).  A group then has exactly one sender and one receiver whose operations match in issue order,
    and traffic of the opposite direction lives on a different communicator, so no ordering between a
    stage's sends and receives can deadlock and nothing has to be issued as a blocking batch.  On CUDA the
    operations are issued from the link's own stream: a send waits (event) for the kernel that produced
    its buffer and the compute stream goes on; a receive is posted ahead of time and the compute stream
    waits for it only when the data is needed (``Work.wait`` is a stream-level wait for NCCL).
    Reference intent: main-pipe.py:79-83,162 (legacy Pipe moves micro-batches on copy streams).
    """

    def __init__(self, group, peer: int, device):
        self.group, self.peer = group, peer
        self.cuda = torch.device(device).type == "cuda"
        self.stream = torch.cuda.Stream(device) if self.cuda else None
        self.pending = []                       # (work, tensor) of sends not joined yet

    def order_after_compute(self):
        """Operations issued from now on start after everything the compute stream has queued so far."""
        if self.cuda:
            self.stream.wait_stream(torch.cuda.current_stream())

    def send(self, t: torch.Tensor):
        if self.cuda:
            ev = torch.cuda.Event()
            ev.record()
            self.stream.wait_event(ev)
            with torch.cuda.stream(self.stream):
                w = dist.isend(t, dst=self.peer, group=self.group)
        else:
            w = dist.isend(t, dst=self.peer, group=self.group)
        self.pending.append((w, t))

    def post_recv(self, buf: torch.Tensor):
        if self.cuda:
            with torch.cuda.stream(self.stream):
                return dist.irecv(buf, src=self.peer, group=self.group)
        return dist.irecv(buf, src=self.peer, group=self.group)

    def join(self):
        """Buffers of finished sends may be reused after this (stream-level on CUDA, blocking on gloo)."""
        for w, _ in self.pending:
            w.wait()
        self.pending = []


class _Slot:
    """Per-micro-batch state of a stage: wire buffers, the autograd endpoints of the live forward, the static
    loss / count tensors of the last stage and, in graph mode, the captured forward / backward graphs."""

    def __init__(self):
        self.act_in = self.out_wire = self.grad_in = self.gin_wire = None
        self.x_in = self.out = None
        self.loss = self.n_valid = self.n_correct = None
        self.fgraph = self.bgraph = None
        self.recompute = False


class PipeStrategy:
    """Trainer strategy for main-pipe.py (dp_size == 1) and main-pipe-ddp.py (dp_size > 1).

    Rank layout for pp x dp: ``rank = pp_rank * dp_size + dp_rank`` - neighbouring stages of one
    replica are ``dp_size`` ranks apart, the replicas of one stage are consecutive ranks.

    Execution: every stage walks its GPipe / 1F1B action list (``schedule_actions``).  Receives of all
    micro-batches are posted up front on the link streams, sends leave from the link streams behind an
    event, so transfers overlap the stage's compute.  On CUDA (native kernels) each micro-batch's forward
    and backward are captured once as CUDA graphs (shapes are static) and replayed - at the default model
    size a micro-batch is a few dozen 5-10 us kernels and the host would otherwise be the bottleneck.
    ``checkpoint`` = never | except_last | always re-runs a micro-batch's forward inside its backward instead
    of keeping its activations (legacy Pipe default ``except_last``, main-pipe.py:79-83).
    """

    static_batches = True      # micro-batches travel between stages in fixed-shape buffers

    def __init__(self, num_stages: int = 0, num_microbatches: int = 0, schedule: str = "gpipe", dp_size: int = 1,
                 checkpoint: str = "never"):
        self.num_stages, self.num_microbatches, self.schedule, self.dp_size = \
            num_stages, num_microbatches, schedule, dp_size
        assert checkpoint in ("never", "except_last", "always")
        self.checkpoint = checkpoint
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
        self.checkpoint = getattr(args, "checkpoint", None) or self.checkpoint
        dp = self.dp_size
        # process groups: every rank must create every group, in the same order
        self.dp_group = None
        if dp > 1:
            for s in range(S):
                g = dist.new_group([s * dp + d for d in range(dp)])
                if s == self.pp_rank:
                    self.dp_group = g
        # one 2-rank group per stage boundary, direction and replica, and one from the last stage to the first
        # (the loss travels there; stage 0 finishes every step last, so it never waits for it)
        self.fwd_in = self.fwd_out = self.bwd_in = self.bwd_out = self.loss_link = None
        for d in range(dp):
            for l in range(S - 1):
                a, b = l * dp + d, (l + 1) * dp + d
                gf, gb = dist.new_group([a, b]), dist.new_group([a, b])
                if rank == a:
                    self.fwd_out, self.bwd_in = _Link(gf, b, self.device), _Link(gb, b, self.device)
                if rank == b:
                    self.fwd_in, self.bwd_out = _Link(gf, a, self.device), _Link(gb, a, self.device)
            if S > 1:
                a, b = d, (S - 1) * dp + d
                gl = dist.new_group([a, b])
                if rank == a:
                    self.loss_link = _Link(gl, b, self.device)
                if rank == b:
                    self.loss_link = _Link(gl, a, self.device)
        self.links = [l for l in (self.fwd_in, self.fwd_out, self.bwd_in, self.bwd_out, self.loss_link) if l is not None]
        self.stage = PipelineStage(model, self.pp_rank, S).to(self.device)
        cuda = self.device.type == "cuda"
        self.native_ok = cuda and _lib.available() and not args.disable_amp
        params = [p for p in self.stage.parameters() if p.requires_grad]
        self.reducer = None
        symm = None
        grad_alloc, holder = None, {}
        transport = "gloo" if not cuda else "nccl"
        if dp > 1:
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
        self.comm_kind = f"p2p:{'nccl' if cuda else 'gloo'}" + (f"+dp:{transport}" if dp > 1 else "")
        self.arena = ParamArena(params, self.device, with_bf16=self.native_ok, with_moments=self.native_ok,
                                grad_alloc=grad_alloc)
        self.optim = FusedAdamW(params, lr=args.learning_rate, device=self.device, arena=self.arena)
        if dp > 1:
            from .ddp import GradReducer
            self.reducer = GradReducer(self.arena, self.dp_group, getattr(args, "bucket_mb", 16.0), transport,
                                       symm, holder.get("buf"), comm_dtype=getattr(args, "comm_dtype", "fp32"))
        self.mb = args.batch_size // self.M
        assert self.mb * self.M == args.batch_size, "batch_size must be divisible by the number of micro-batches"
        self.T = args.sequence_length - 1
        self.wire = torch.bfloat16 if (cuda and not args.disable_amp) else torch.float32
        self.has_loss = self.stage.is_last
        self.use_graph = self.native_ok and not getattr(args, "disable_cuda_graph", False)
        self.graph_warmup = 2                  # eager steps before the per-micro-batch graphs are captured
        self.steps_done = 0
        self.slots: List[_Slot] = []
        self._static = {}
        self._pool = None
        self._cap_stream = None
        self._warm_links()

    def _warm_links(self):
        """Create the link communicators now, boundary by boundary from the first stage on (a chain, so the lazy
        NCCL initialisation of one link never waits on a rank that is itself waiting on an earlier link)."""
        one = torch.zeros(1, dtype=torch.float32, device=self.device)
        for link, sending in ((self.fwd_in, False), (self.bwd_out, True), (self.fwd_out, True), (self.bwd_in, False)):
            if link is None:
                continue
            (dist.send if sending else dist.recv)(one, link.peer, group=link.group)
        if self.loss_link is not None:
            (dist.send if self.stage.is_last else dist.recv)(one, self.loss_link.peer, group=self.loss_link.group)
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)

    def train(self):
        self.stage.train()

    def eval(self):
        self.stage.eval()

    def _autocast(self, amp):
        if amp and self.device.type == "cuda":
            return torch.autocast(device_type="cuda", dtype=torch.bfloat16)
        return nullcontext()

    # ------------------------------------------------------------------------------
    def _make_slots(self, dim: int):
        st, dev = self.stage, self.device
        shape = (self.mb, self.T, dim)
        for k in range(self.M):
            sl = _Slot()
            if not st.is_first:
                sl.act_in = torch.zeros(shape, dtype=self.wire, device=dev)
                sl.gin_wire = torch.zeros(shape, dtype=self.wire, device=dev)
            if not st.is_last:
                sl.out_wire = torch.zeros(shape, dtype=self.wire, device=dev)
                sl.grad_in = torch.zeros(shape, dtype=self.wire, device=dev)
            sl.recompute = self.checkpoint == "always" or (self.checkpoint == "except_last" and k < self.M - 1)
            self.slots.append(sl)

    def _micro(self, t, i):
        return None if t is None else t[i * self.mb:(i + 1) * self.mb]

    def _run_forward(self, sl: _Slot, i: int, batch, targets, amp: bool, grad: bool):
        """Forward of micro-batch i on this stage.  grad=False: nothing is kept for a backward pass."""
        st = self.stage
        with self._autocast(amp), (nullcontext() if grad else torch.no_grad()):
            if st.is_first:
                x_in = None
                x = st.forward_hidden(None, self._micro(batch["input_ids"], i), self._micro(batch["position_ids"], i),
                                      self._micro(batch.get("mask"), i))
            else:
                x_in = sl.act_in.float()
                if x_in is sl.act_in:             # fp32 wire: the autograd leaf must not be the receive buffer itself
                    x_in = x_in.clone()
                if grad:
                    x_in.requires_grad_(True)
                x = st.forward_hidden(x_in, mask=self._micro(batch.get("mask"), i))
            if st.is_last:
                loss, n_valid, n_correct = st.loss(x, self._micro(targets, i), grad_scale=1.0 / self.M)
                out = loss
                for name, v in (("loss", loss.detach()), ("n_valid", n_valid), ("n_correct", n_correct)):
                    v = v.float() if torch.is_tensor(v) else torch.tensor(float(v), device=self.device)
                    cur = getattr(sl, name)
                    if cur is None:
                        setattr(sl, name, v.clone())
                    else:
                        cur.copy_(v)
            else:
                out = x
                sl.out_wire.copy_(x.detach())
        if grad:
            sl.x_in, sl.out = x_in, out
        return out

    def _run_backward(self, sl: _Slot, i: int, batch, targets, amp: bool):
        st = self.stage
        if sl.recompute:                       # activation checkpointing: redo the forward, now with autograd
            self._run_forward(sl, i, batch, targets, amp, grad=True)
        if st.is_last:
            (sl.out * (1.0 / self.M)).backward()
        else:
            torch.autograd.backward(sl.out, sl.grad_in.to(sl.out.dtype))
        if not st.is_first:
            sl.gin_wire.copy_(sl.x_in.grad)
        sl.x_in = sl.out = None

    def _capture(self, fn):
        """Capture ``fn`` into a CUDA graph (shared memory pool, no device-wide synchronisation: transfers of
        other micro-batches are in flight on the link streams while a stage captures)."""
        if self._cap_stream is None:
            self._cap_stream = torch.cuda.Stream(self.device)
            self._pool = torch.cuda.graph_pool_handle()
        g = torch.cuda.CUDAGraph()
        cur = torch.cuda.current_stream()
        self._cap_stream.wait_stream(cur)
        with torch.cuda.stream(self._cap_stream):
            g.capture_begin(pool=self._pool, capture_error_mode="thread_local")
            try:
                fn()
            finally:
                g.capture_end()
        cur.wait_stream(self._cap_stream)
        return g

    def _forward_action(self, sl, i, batch, targets, amp, graphed):
        grad = not sl.recompute
        if not graphed:
            self._run_forward(sl, i, batch, targets, amp, grad)
            return
        if sl.fgraph is None:
            sl.fgraph = self._capture(lambda: self._run_forward(sl, i, batch, targets, amp, grad))
        sl.fgraph.replay()

    def _backward_action(self, sl, i, batch, targets, amp, graphed):
        if not graphed:
            self._run_backward(sl, i, batch, targets, amp)
            return
        if sl.bgraph is None:
            sl.bgraph = self._capture(lambda: self._run_backward(sl, i, batch, targets, amp))
        sl.bgraph.replay()

    def _static_inputs(self, batch, targets):
        """Graph mode: the captured kernels read the batch from fixed addresses."""
        out = {}
        for name, t in (("input_ids", batch["input_ids"]), ("position_ids", batch["position_ids"]), ("targets", targets)):
            buf = self._static.get(name)
            if buf is None or buf.shape != t.shape:
                buf = self._static[name] = torch.empty_like(t, device=self.device)
            buf.copy_(t, non_blocking=True)
            out[name] = buf
        return {"input_ids": out["input_ids"], "position_ids": out["position_ids"]}, out["targets"]

    def train_step(self, batch, targets, amp=True):
        from ..utils import nvtx
        M, S, st = self.M, self.S, self.stage
        graphed = self.use_graph and amp and self.steps_done >= self.graph_warmup
        capturing = graphed and (not self.slots or self.slots[0].fgraph is None)
        if not self.slots:
            self._make_slots(st.dim)
        if graphed:
            batch, targets = self._static_inputs(batch, targets)   # the key-padding mask is a no-op for the loss: dropped
        self.optim.zero_grad()
        if self.reducer is not None:
            self.reducer.begin()
            self.reducer.active = False              # grads accumulate over micro-batches: reduce at the end
        for l in self.links:
            l.order_after_compute()
        # Receives are posted ahead of the compute that needs them - but only once every kernel of a step has run
        # at least once.  A posted NCCL receive is a kernel spinning on the GPU, and the FIRST launch of any kernel
        # may have to load its module, which waits for running kernels: a gradient receive posted at the start of
        # a step can only finish after this very stage has launched its forward kernels, so during the first steps
        # (and in the step that captures the graphs) receives are posted one at a time, right where they are needed.
        acts = grads = None
        settled = self.steps_done >= self.graph_warmup + 2
        if settled and not capturing:
            acts = [self.fwd_in.post_recv(sl.act_in) for sl in self.slots] if not st.is_first else None
            grads = [self.bwd_in.post_recv(sl.grad_in) for sl in self.slots] if not st.is_last else None
        for kind, i in schedule_actions(self.schedule, self.pp_rank, S, M):
            sl = self.slots[i]
            if kind == "F":
                if not st.is_first:
                    (acts[i] if acts is not None else self.fwd_in.post_recv(sl.act_in)).wait()
                with nvtx.phase(f"forward mb{i}"):
                    self._forward_action(sl, i, batch, targets, amp, graphed)
                if not st.is_last:
                    self.fwd_out.send(sl.out_wire)
            else:
                if not st.is_last:
                    (grads[i] if grads is not None else self.bwd_in.post_recv(sl.grad_in)).wait()
                with nvtx.phase(f"backward mb{i}"):
                    self._backward_action(sl, i, batch, targets, amp, graphed)
                if not st.is_first:
                    self.bwd_out.send(sl.gin_wire)
        with nvtx.phase("grad-sync + optimizer"):
            if self.reducer is not None:
                self.reducer.active = True
                self.reducer.finish()
            self.optim.step()
        self.steps_done += 1
        # the loss goes from the last stage to the first (rank 0 prints); other stages report 0
        loss = torch.zeros((), dtype=torch.float32, device=self.device)
        if st.is_last:
            loss = torch.stack([sl.loss for sl in self.slots]).mean()
            if self.loss_link is not None:
                self._loss_out = loss.reshape(1).clone()
                self.loss_link.send(self._loss_out)
        elif st.is_first and self.loss_link is not None:
            if getattr(self, "_loss_in", None) is None:
                self._loss_in = torch.zeros(1, dtype=torch.float32, device=self.device)   # (a fresh zeros() each step would
            self.loss_link.post_recv(self._loss_in).wait()                                # race its fill kernel with the receive)
            loss = self._loss_in[0].clone()
        for l in self.links:
            l.join()
        if st.is_first and self.dp_size > 1:
            loss = loss.clone()
            dist.all_reduce(loss, group=self.dp_group)
            loss = loss / self.dp_size
        return loss.detach()

    @torch.no_grad()
    def eval_step(self, batch, targets, amp=True):
        st = self.stage
        if not self.slots:
            self._make_slots(st.dim)
        for l in self.links:
            l.order_after_compute()
        for i in range(self.M):
            sl = self.slots[i]
            if not st.is_first:
                self.fwd_in.post_recv(sl.act_in).wait()
            self._run_forward(sl, i, batch, targets, amp, grad=False)
            if not st.is_last:
                self.fwd_out.send(sl.out_wire)
        for l in self.links:
            l.join()
        res = torch.zeros(3, device=self.device)
        if st.is_last:
            res[0] = torch.stack([sl.loss for sl in self.slots]).mean()
            res[1] = torch.stack([sl.n_valid for sl in self.slots]).sum()
            res[2] = torch.stack([sl.n_correct for sl in self.slots]).sum()
        dist.all_reduce(res)                      # only last-stage ranks contribute
        res = res / self.dp_size if self.dp_size > 1 else res
        return res[0], res[1], res[2]

    def all_reduce_mean(self, t):
        return t                                   # eval_step already returned group-wide values

    def close(self):
        if self.reducer is not None:
            self.reducer.close()
        self.slots = []

    def optimizer_state(self):
        return self.optim.state_dict()

    def load_optimizer_state(self, sd):
        self.optim.load_state_dict(sd)

    # ------------------------------------------------------------------------------
    def gather_full_model(self):
        """Collective: rank 0 gets a complete, up-to-date copy of the model (its own ``TransformerDecoderLM``
        object with every stage's weights loaded), the other ranks ``None``.  Used for the per-epoch sample
        generation of the reference (main-pipe.py:210-214), which needs all stages in one place."""
        sd = self.state_dict()
        if dist.get_rank() != 0:
            return None
        if getattr(self, "_gen_model", None) is None:
            import copy
            full = self.stage._full
            self._gen_model = copy.deepcopy(full).to("cpu").float().to(self.device)
        self._gen_model.load_state_dict(sd)
        for p in self._gen_model.parameters():
            if hasattr(p, "_b200_bf16_version"):
                p._b200_bf16_version = -1
        return self._gen_model

    def generation_model(self):
        return self.gather_full_model()

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
