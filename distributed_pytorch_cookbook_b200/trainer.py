"""The one shared training / validation / generation / checkpoint loop.

The reference copy-pastes this loop into each recipe (main-single.py:80-151,
main-ddp.py:105-187, main-fsdp.py:119-202, main-pipe.py:150-221).  Here every recipe is
``parse flags -> pick a parallel strategy -> Trainer(args, strategy).run()``.

Behaviour kept from the reference: per-rank ``--batch_size``; tqdm description strings
``"[training] Epoch e/E | loss: x.xxx"`` refreshed every PRINT_FREQ=8 steps (rank 0 only);
validation loss + token accuracy (%) with running means; three fixed generation prompts
after every epoch; one checkpoint at the end under ``checkpoints/checkpoint-<time>.pt``.
Deliberate fixes: the sampler gets ``set_epoch``; no host sync per step (the loss stays on
the device and is read back once per PRINT_FREQ window); validation metrics are reduced
with one fused 2-float all-reduce per batch instead of two.
"""
from __future__ import annotations

import time
from contextlib import nullcontext
from typing import Optional

import torch
import torch.distributed as dist

from .data import (PinnedPrefetcher, ShardedSampler, SyntheticTokenDataset, SyntheticTokenizer,
                   get_dataset, get_tokenizer, make_loader, transform_dataset)
from .models import TransformerDecoderLM
from .utils import (JsonlLogger, StepTimer, generate, load_checkpoint, prepare_batch,
                    save_checkpoint, seed_everything)
from .utils import nvtx

PRINT_FREQ = 8
PROMPTS = ("The big brown cat ", "One day, ", "She said ")


def _tqdm(it, disable=False):
    try:
        from tqdm import tqdm
        return tqdm(it, disable=disable)
    except Exception:  # pragma: no cover
        class _PB:
            def __init__(self, it):
                self.it = it

            def __iter__(self):
                return iter(self.it)

            def set_description(self, s):
                pass
        return _PB(it)


def build_model(args, vocab_size: int) -> TransformerDecoderLM:
    """Model construction shared by all recipes (reference main-single.py:26-33)."""
    return TransformerDecoderLM(dim=args.dim, head_dim=args.head_dim, heads=args.heads,
                                num_layers=args.num_layers, vocab_size=vocab_size,
                                max_position_embeddings=args.sequence_length)


def check_device_error(where: str = ""):
    """Raise if a kernel reported a failure through the device error word (a bounded wait that timed out: a
    cross-GPU barrier whose peer never arrived, a GEMM pipeline stage that never filled).  Such a kernel has
    returned early - the collective or tile it stood for did not happen - so continuing would train on
    unreduced gradients.  The condition is fatal: the caller should let the exception end the process
    (torchrun then tears the other ranks down)."""
    from .ops import _lib
    if not (_lib.loaded() and torch.cuda.is_available() and _lib.has("b200_device_error")):
        return
    import ctypes
    code = ctypes.c_int(0)
    _lib.call("b200_device_error", ctypes.addressof(code), 0)
    if code.value != 0:
        kinds = {101: "GEMM producer wait", 102: "GEMM MMA wait", 103: "GEMM epilogue wait", 104: "GEMM smem alignment",
                 105: "fused all-gather wait", 301: "cross-GPU signal", 302: "cross-GPU wait (a peer never arrived)"}
        raise RuntimeError(f"device error {code.value} ({kinds.get(code.value, 'kernel-side wait timed out')}) "
                           f"detected {where}: a kernel gave up waiting and skipped its work; aborting the run")


def build_data(args, rank: int = 0, world: int = 1, data_parallel: bool = False, static_batches: bool = False):
    """tokenizer + train/val datasets + loaders (+ samplers when data-parallel)."""
    synthetic = getattr(args, "synthetic", False)
    if synthetic:
        tokenizer = SyntheticTokenizer(getattr(args, "vocab_size", 50257), args.sequence_length)
        n = getattr(args, "synthetic_samples", 4096)
        train = SyntheticTokenDataset(n, args.sequence_length, tokenizer.vocab_size, seed=1234)
        val = SyntheticTokenDataset(max(n // 8, args.batch_size * world), args.sequence_length,
                                    tokenizer.vocab_size, seed=4321)
    else:
        tokenizer = get_tokenizer()
        train, val = get_dataset(slice_size=args.dataset_slice)
        train = transform_dataset(train, tokenizer, max_length=args.sequence_length, num_proc=args.num_workers)
        val = transform_dataset(val, tokenizer, max_length=args.sequence_length, num_proc=args.num_workers)
    tokenizer.pad_token_id = 2  # reference main-single.py:23
    tsamp = vsamp = None
    if data_parallel and world > 1:
        tsamp = ShardedSampler(len(train), world, rank, shuffle=True, seed=getattr(args, "seed", 0))
        vsamp = ShardedSampler(len(val), world, rank, shuffle=False)
    workers = 0 if synthetic else args.num_workers
    train_loader = make_loader(train, args.batch_size, sampler=tsamp, shuffle=tsamp is None,
                               num_workers=workers, drop_last=True)
    # the reference's validation loader keeps the last partial batch (main-single.py:65-71); only engines that move
    # fixed-shape micro-batches between stages (pipeline) need full batches
    val_loader = make_loader(val, args.batch_size, sampler=vsamp, shuffle=False,
                             num_workers=workers, drop_last=static_batches)
    return tokenizer, train_loader, val_loader, tsamp


class Trainer:
    def __init__(self, args, strategy, rank: int = 0, world: int = 1, device="cpu",
                 data_parallel: bool = False, is_main: Optional[bool] = None):
        """``rank``/``world`` describe the *data-parallel* position (which shard of the data this
        process reads); ``is_main`` who prints and saves (default: global rank 0)."""
        self.args, self.strategy = args, strategy
        self.rank, self.world = rank, world
        self.device = torch.device(device)
        self.data_parallel = data_parallel
        if is_main is None:
            is_main = (dist.get_rank() == 0) if dist.is_initialized() else rank == 0
        self.main = is_main
        self.metrics = JsonlLogger(getattr(args, "metrics_jsonl", None), enabled=self.main)

    # ------------------------------------------------------------------------------
    def run(self):
        args = self.args
        seed_everything(getattr(args, "seed", 0))
        nvtx.enable(getattr(args, "profile", False))
        if nvtx.enabled():
            args.disable_cuda_graph = True      # phases are only visible when the step is not one graph launch
        tokenizer, train_loader, val_loader, tsamp = build_data(
            args, self.rank, self.world, self.data_parallel,
            static_batches=getattr(self.strategy, "static_batches", False))
        self.tokenizer = tokenizer
        model = build_model(args, tokenizer.vocab_size)
        resume_extra = None
        if getattr(args, "resume", None):
            _, resume_extra = load_checkpoint(model, args.resume, rank=self._global_rank())
        self.strategy.setup(model, args, self.device)
        start_epoch = 0
        if resume_extra is not None:
            start_epoch = self._restore(resume_extra)
        amp = not args.disable_amp
        timer = StepTimer(self.device)
        tokens_per_step = args.batch_size * (args.sequence_length - 1) * (self.world if self.data_parallel else 1)

        for ei in range(start_epoch, args.epochs):
            if tsamp is not None:
                tsamp.set_epoch(ei)
            pb = _tqdm(PinnedPrefetcher(train_loader, self.device), disable=not self.main)
            pb.set_description(f"[training] Epoch {ei+1}/{args.epochs} | loss: ?????")
            self.strategy.train()
            window = []
            timer.start()
            for i, batch in enumerate(pb):
                if getattr(args, "max_steps", 0) and i >= args.max_steps:
                    break
                batch, targets = prepare_batch(batch, tokenizer.pad_token_id, self.device)
                with nvtx.phase("train_step"):
                    loss = self.strategy.train_step(batch, targets, amp=amp)
                window.append(loss)
                if i > 0 and not i % PRINT_FREQ:
                    ms = timer.stop_ms() / len(window)
                    mean = float(torch.stack([w.detach().float() for w in window]).mean())
                    check_device_error(f"at epoch {ei + 1} step {i}")     # the loss read-back above already synchronised
                    pb.set_description(f"[training] Epoch {ei+1}/{args.epochs} | loss: {mean:.3f}")
                    self.metrics.log(phase="train", epoch=ei, step=i, loss=mean, ms_per_step=ms,
                                     tokens_per_s=tokens_per_step / ms * 1e3)
                    window = []
                    timer.start()

            self.validate(ei, val_loader, amp=True)  # reference validates under autocast always (:119)
            check_device_error(f"after epoch {ei + 1}")
            if not getattr(args, "no_generate", False):
                self.sample(tokenizer)
            if dist.is_initialized():
                dist.barrier()
            self.epochs_done = ei + 1

        path = None
        if not getattr(args, "no_save", False):
            if dist.is_initialized():
                dist.barrier()
            check_device_error("before saving the checkpoint")
            state = self.strategy.state_dict()  # collective for FSDP / pipeline
            if self.main:
                path = save_checkpoint(state, getattr(args, "checkpoint_dir", "checkpoints"))
            self._save_resume_state(path)
        self.metrics.close()
        return path

    # ------------------------------------------------------------------------------
    def _global_rank(self) -> int:
        return dist.get_rank() if dist.is_initialized() else 0

    def _save_resume_state(self, path):
        """Optimizer moments + step counter + epochs done, one ``<checkpoint>.resume.rank<r>`` per rank (FSDP
        shards and pipeline stages own different slices of the optimizer state).  The weights file itself stays
        the plain state_dict the reference loads."""
        from pathlib import Path
        if not hasattr(self.strategy, "optimizer_state"):
            return
        name = [str(path) if path is not None else None]
        if dist.is_initialized():
            dist.broadcast_object_list(name, src=0)
        if name[0] is None:
            return
        extra = {"optimizer": self.strategy.optimizer_state(), "epochs_done": getattr(self, "epochs_done", 0),
                 "world": dist.get_world_size() if dist.is_initialized() else 1}
        torch.save(extra, Path(name[0] + f".resume.rank{self._global_rank()}"))

    def _restore(self, extra) -> int:
        world = dist.get_world_size() if dist.is_initialized() else 1
        if extra.get("world") != world or not hasattr(self.strategy, "load_optimizer_state"):
            if self.main:
                print("[resume] optimizer state was saved for a different layout; continuing from the weights only",
                      flush=True)
            return 0
        self.strategy.load_optimizer_state(extra["optimizer"])
        if self.main:
            print(f"[resume] optimizer state restored ({extra.get('epochs_done', 0)} epochs done)", flush=True)
        return int(extra.get("epochs_done", 0)) if getattr(self.args, "resume_epochs", False) else 0

    # ------------------------------------------------------------------------------
    @torch.no_grad()
    def validate(self, ei: int, val_loader, amp: bool = True):
        args = self.args
        self.strategy.eval()
        pb = _tqdm(PinnedPrefetcher(val_loader, self.device), disable=not self.main)
        pb.set_description(f"[validation] Epoch {ei+1}/{args.epochs} | loss: ?????, accuracy: ?????")
        total = torch.zeros(2, device=self.device)
        n = 0
        for i, batch in enumerate(pb):
            if getattr(args, "max_steps", 0) and i >= args.max_steps:
                break
            batch, targets = prepare_batch(batch, self.tokenizer.pad_token_id, self.device)
            out = self.strategy.eval_step(batch, targets, amp=amp)
            if out is None:        # pipeline ranks that do not hold the loss
                continue
            loss, n_valid, n_correct = out
            pair = torch.stack([loss.float(), n_correct.float() / n_valid.clamp(min=1).float() * 100.0])
            if self.data_parallel and self.world > 1:
                self.strategy.all_reduce_mean(pair)
            total += pair
            n += 1
            if self.main and (i % PRINT_FREQ == 0):
                t = (total / n).tolist()
                pb.set_description(f"[validation] Epoch {ei+1}/{args.epochs} | loss: {t[0]:.3f}, accuracy: {t[1]:.2f}")
        if n:
            t = (total / n).tolist()
            self.metrics.log(phase="val", epoch=ei, loss=t[0], accuracy=t[1])
            if self.main:
                print(f"[validation] Epoch {ei+1}/{args.epochs} | loss: {t[0]:.3f}, accuracy: {t[1]:.2f}", flush=True)
            return t
        return None

    def sample(self, tokenizer):
        gen_model = self.strategy.generation_model()
        if gen_model is None or not self.main:
            return
        gen_model.eval()
        print("Argmax sampling from model")
        for p in PROMPTS:
            print(generate(gen_model, p, tokenizer, self.device, use_cache=True))   # same tokens as the reference's full re-run


class SingleStrategy:
    """One device (CUDA or CPU): reference main-single.py.  On CUDA with AMP the whole
    step (forward, backward, fused AdamW) is captured once into a CUDA graph and replayed."""

    def __init__(self):
        self.model = None

    def setup(self, model, args, device):
        from .optim import build_optimizer
        from .runtime.graph import GraphedStep
        self.device = torch.device(device)
        self.model = model.to(self.device)
        self.optim = build_optimizer(self.model, lr=args.learning_rate, device=self.device)
        self.use_graph = (self.device.type == "cuda" and not args.disable_amp
                          and not getattr(args, "disable_cuda_graph", False))
        self._graphed: Optional[GraphedStep] = None

    def train(self):
        self.model.train()

    def eval(self):
        self.model.eval()

    def _raw_step(self, input_ids, position_ids, targets, mask, amp):
        ctx = torch.autocast(device_type="cuda", dtype=torch.bfloat16) if (amp and self.device.type == "cuda") else nullcontext()
        self.optim.zero_grad()
        with ctx, nvtx.phase("forward"):
            loss, _, _ = self.model.loss(input_ids, position_ids, targets, mask=mask)
        with nvtx.phase("backward"):
            loss.backward()
        with nvtx.phase("optimizer"):
            self.optim.step()
        return loss.detach()

    def optimizer_state(self):
        return self.optim.state_dict()

    def load_optimizer_state(self, sd):
        self.optim.load_state_dict(sd)

    def train_step(self, batch, targets, amp=True):
        if self.use_graph:
            from .runtime.graph import GraphedStep
            if self._graphed is None:
                self._graphed = GraphedStep(lambda i, p, t: self._raw_step(i, p, t, None, True))
            # the key-padding mask is a no-op for the loss and is dropped here
            return self._graphed(batch["input_ids"], batch["position_ids"], targets)
        return self._raw_step(batch["input_ids"], batch["position_ids"], targets, batch.get("mask"), amp)

    @torch.no_grad()
    def eval_step(self, batch, targets, amp=True):
        ctx = torch.autocast(device_type="cuda", dtype=torch.bfloat16) if (amp and self.device.type == "cuda") else nullcontext()
        with ctx:
            return self.model.loss(batch["input_ids"], batch["position_ids"], targets, mask=batch.get("mask"))

    def all_reduce_mean(self, t):
        return t

    def launches_per_step(self):
        """native entry-point calls captured in one CUDA-graph replay (None before capture)."""
        return self._graphed.native_calls_per_replay if (self._graphed and self._graphed.graph) else None

    def generation_model(self):
        return self.model

    def state_dict(self):
        from .utils import canonical_state_dict
        return canonical_state_dict(self.model)
