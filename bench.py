#!/usr/bin/env python
"""Headline benchmark: GPT training tokens/sec (BASELINE.json), device-timed, max over ranks.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Config = BASELINE.md row 2: the reference's default GPT (dim 256, head_dim 32, heads 8,
layers 8, seq 256 -> 255 tokens/sample, vocab 50257), bf16, batch 64 per rank, AdamW step
included; N = 1 is the single-GPU recipe, N > 1 the DDP recipe (weak scaling) launched by
``python -m torch.distributed.run``.  Synthetic tokens, random-init weights.

Two numbers per run:
  value  - K steps replayed back to back from device-resident batches (kernel-only view);
  e2e    - K steps through the public API (Strategy.train_step fed by utils.prepare_batch)
           where every step copies its batch from pinned host memory and reads the loss back.
Both are bracketed by barrier + cuda synchronize, timed with CUDA events, max over ranks.
``--impl reference`` runs the reference's own model / loop from baseline/_ref instead.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path
from types import SimpleNamespace

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

CFG = dict(dim=256, head_dim=32, heads=8, num_layers=8, sequence_length=256, vocab_size=50257,
           batch_size=64, learning_rate=1e-4)
METRIC = "GPT training tokens/sec (device-timed, max over ranks)"


# ------------------------------------------------------------------------------------------
def world():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None
        return self

    def __exit__(self, *exc):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
            try:
                out, _ = self.proc.communicate(timeout=5)
            except Exception:
                self.proc.kill()
                out = ""
            self.lines = [l for l in out.splitlines() if l.strip()]

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except Exception:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "samples": len(sm)}


def timed_blocks(fn, steps: int, device, sampler=None, min_total_s: float = 1.0, max_blocks: int = 40):
    """The K-step timed region, repeated: every block is ``timed`` (barrier + synchronize on both sides, CUDA
    events, max over ranks) over exactly K steps; blocks are added until at least ``min_total_s`` of device time
    has been measured.  Returns the per-block times in ms.  One clock sampler spans all blocks."""
    out = []
    ctx = sampler if sampler is not None else _Null()
    with ctx:
        while True:
            out.append(timed(fn, steps, device))
            if sum(out) >= min_total_s * 1e3 or len(out) >= max_blocks:
                break
    return out


def timed(fn, steps: int, device, sampler=None):
    """barrier + sync, K x fn under CUDA events, sync + barrier; returns ms (max over ranks)."""
    if dist.is_initialized():
        dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx = sampler if sampler is not None else _Null()
    with ctx:
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        torch.cuda.synchronize(device)
    if dist.is_initialized():
        dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=device)
    if dist.is_initialized():
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms)


class _Null:
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False


def host_batches(n: int, seed: int):
    """n pinned host batches in the reference's dataset format ({input_ids, attention_mask} int64 [B,S])."""
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n):
        ids = torch.randint(0, CFG["vocab_size"], (CFG["batch_size"], CFG["sequence_length"]), generator=g)
        out.append({"input_ids": ids.pin_memory(), "attention_mask": torch.ones_like(ids).pin_memory()})
    return out


def result_line(impl, value, ms, e2e_value, n, steps, warmup, clocks, launches, details, h2d, d2h):
    """``config`` holds only what both arms share (model / batch / sequence / parallelism); everything that is
    specific to an arm (graph capture, transport, compile mode, losses, block statistics) goes to ``details``."""
    baseline = None  # BASELINE.md publishes no number ("none published")
    line = {
        "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": n, "steps": steps, "warmup": warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": (value / baseline) if baseline else None, "dtype": "bf16", "data": "synthetic",
        "impl": impl,
        "config": {"model": "GPT dim256 head_dim32 heads8 layers8 vocab50257 (reference default, 32.1M params)",
                   "global_batch": CFG["batch_size"] * n, "seq_len": CFG["sequence_length"],
                   "tokens_per_sample": CFG["sequence_length"] - 1,
                   "parallelism": "single" if n == 1 else f"dp{n}",
                   "optimizer": "AdamW lr1e-4 (step inside timed region)",
                   "l2": "no flush: per-step working set (0.5 GB params+optimizer state, >1 GB activations) "
                         "exceeds the 126 MB L2"},
        "details": details,
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches,
    }
    return line


# ------------------------------------------------------------------------------------------
def run_ours(args):
    from distributed_pytorch_cookbook_b200.ops import _lib
    from distributed_pytorch_cookbook_b200.trainer import SingleStrategy, build_model
    from distributed_pytorch_cookbook_b200.utils import prepare_batch, seed_everything

    rank, local_rank, n = world()
    device = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(device)
    if n > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=device)
    assert _lib.available(), f"native sm_100a library not loaded: {_lib.load_error()}"
    seed_everything(0)
    ns = SimpleNamespace(**CFG, disable_amp=False, disable_cuda_graph=False, bucket_mb=args.bucket_mb,
                         comm=args.comm, epochs=1, comm_dtype=args.comm_dtype)
    model = build_model(ns, CFG["vocab_size"])
    if n == 1:
        strategy = SingleStrategy()
    else:
        from distributed_pytorch_cookbook_b200.parallel.ddp import DDPStrategy
        strategy = DDPStrategy()
    strategy.setup(model, ns, device)
    strategy.train()

    batches = host_batches(max(4, min(args.steps, 16)), seed=100 + rank)
    pad = 2

    def e2e_step(i):
        b, t = prepare_batch(batches[i % len(batches)], pad, device)   # H2D from pinned memory
        loss = strategy.train_step(b, t, amp=True)
        return loss

    # warm-up (also triggers CUDA-graph capture inside the strategy)
    for i in range(max(args.warmup, 5)):
        e2e_step(i)
    torch.cuda.synchronize(device)

    # kernel-only view: device-resident batches, no host traffic in the loop
    dev_batches = [prepare_batch(b, pad, device) for b in batches]
    torch.cuda.synchronize(device)
    launches0 = _lib.launch_count()
    per_step_launches = strategy.launches_per_step() if hasattr(strategy, "launches_per_step") else None

    def dev_step(i):
        b, t = dev_batches[i % len(dev_batches)]
        strategy.train_step(b, t, amp=True)

    sampler = ClockSampler(local_rank) if rank == 0 else None
    blocks = timed_blocks(dev_step, args.steps, device, sampler, args.min_seconds)
    ms_total = statistics.median(blocks)
    clocks = sampler.summary() if sampler is not None else None
    launches = (_lib.launch_count() - launches0) // len(blocks)
    if per_step_launches:
        launches = per_step_launches * args.steps

    host_losses = []
    # D2H read of every step's loss, the way a training loop that logs does it without stalling the
    # GPU: an async copy into pinned memory right behind the step, read on the host one step later
    # (the last one before the timed region closes).  K steps -> K values read inside the region.
    loss_host = [torch.empty((), dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event() for _ in range(2)]

    def e2e_timed(i):
        loss = e2e_step(i)
        loss_host[i & 1].copy_(loss.detach().float(), non_blocking=True)
        loss_ev[i & 1].record()
        if i > 0:
            loss_ev[(i - 1) & 1].synchronize()
            host_losses.append(float(loss_host[(i - 1) & 1]))
        if i == args.steps - 1:
            loss_ev[i & 1].synchronize()
            host_losses.append(float(loss_host[i & 1]))

    e2e_blocks = timed_blocks(e2e_timed, args.steps, device, None, args.min_seconds)
    ms_e2e = statistics.median(e2e_blocks)
    assert len(host_losses) == args.steps * len(e2e_blocks)

    # exposed communication per step = step time minus the same step with the all-reduces skipped
    exposed_comm_ms = None
    if n > 1 and hasattr(strategy, "reducer"):
        try:
            strategy.reducer.skip_comm = True
            strategy._graphed = None                      # re-capture without the collective kernels
            for i in range(6):
                dev_step(i)
            torch.cuda.synchronize(device)
            ms_nocomm = statistics.median(timed_blocks(dev_step, args.steps, device, None, args.min_seconds / 2))
            exposed_comm_ms = max(ms_total - ms_nocomm, 0.0) / args.steps
        except Exception as e:  # noqa: BLE001  (measurement extra: never lose the headline line)
            print(f"[bench] exposed-communication measurement failed: {e}", file=sys.stderr, flush=True)
    tokens_per_step = CFG["batch_size"] * (CFG["sequence_length"] - 1) * n
    value = tokens_per_step * args.steps / ms_total * 1e3
    e2e_value = tokens_per_step * args.steps / ms_e2e * 1e3
    h2d = sum(v.numel() * v.element_size() for v in prepare_batch(batches[0], pad, "cpu")[0].values()) + \
        CFG["batch_size"] * (CFG["sequence_length"] - 1) * 8
    details = {"cuda_graph": bool(getattr(strategy, "use_graph", False)),
               "final_loss": host_losses[-1] if host_losses else None,
               "comm": getattr(strategy, "comm_kind", None),
               "exposed_comm_ms_per_step": exposed_comm_ms,
               "timed_blocks": len(blocks), "timed_seconds": sum(blocks) / 1e3,
               "ms_per_step_min": min(blocks) / args.steps, "ms_per_step_max": max(blocks) / args.steps,
               "e2e_ms_per_step": ms_e2e / args.steps}
    if hasattr(strategy, "close"):
        strategy.close()
    del strategy, model, dev_batches
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    line = result_line("ours", value, ms_total / args.steps, e2e_value, n, args.steps, args.warmup,
                       clocks, launches, details, h2d, 4)
    if n > 1 and not args.no_selfcheck:
        line["selfcheck"] = selfcheck(device)
    if n > 1 and not args.no_recipes:
        line["recipes"] = run_recipes("ours", n, args)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


def selfcheck(device):
    """Multi-GPU correctness the 1-GPU test tier cannot see: the same 3 training steps (default GPT, fixed seed,
    one batch per rank) through (a) DDP on the own NVLink all-reduce, (b) DDP on NCCL, (c) FSDP on the own
    reduce-scatter / all-gather kernels.  All three must report the same losses."""
    from distributed_pytorch_cookbook_b200.ops import _lib
    from distributed_pytorch_cookbook_b200.parallel.ddp import DDPStrategy
    from distributed_pytorch_cookbook_b200.parallel.fsdp import FSDPStrategy
    from distributed_pytorch_cookbook_b200.trainer import build_model
    from distributed_pytorch_cookbook_b200.utils import prepare_batch, seed_everything
    import ctypes
    rank = dist.get_rank()
    out = {}
    try:
        batches = host_batches(3, seed=500 + rank)
        runs = {}
        for name, make, comm in (("ddp_symm", DDPStrategy, "symm"), ("ddp_nccl", DDPStrategy, "nccl"),
                                 ("fsdp_symm", FSDPStrategy, "symm")):
            seed_everything(0)
            ns = SimpleNamespace(**CFG, disable_amp=False, disable_cuda_graph=False, bucket_mb=16.0, comm=comm,
                                 epochs=1, cpu_offload=False)
            st = make()
            st.setup(build_model(ns, CFG["vocab_size"]), ns, device)
            st.train()
            losses = []
            for b in batches:
                bb, t = prepare_batch(b, 2, device)
                losses.append(st.train_step(bb, t, amp=True).float().reshape(1).clone())
            ls = torch.cat(losses)
            dist.all_reduce(ls)
            runs[name] = (ls / dist.get_world_size()).tolist()
            st.close()
            del st
            torch.cuda.synchronize(device)
        err = ctypes.c_int(0)
        _lib.call("b200_device_error", ctypes.addressof(err), 0)
        ref = runs["ddp_nccl"]
        out = {"losses": runs, "device_error": err.value,
               "max_abs_diff_vs_nccl": {k: max(abs(a - b) for a, b in zip(v, ref)) for k, v in runs.items()}}
        out["ok"] = bool(err.value == 0 and all(d < 2e-2 for d in out["max_abs_diff_vs_nccl"].values())
                         and all(x == x for v in runs.values() for x in v))
    except Exception as e:  # noqa: BLE001
        out = {"ok": False, "error": repr(e)[:300]}
    import gc
    gc.collect()
    torch.cuda.empty_cache()
    return out


def run_recipes(impl, n, args):
    """BASELINE.json configs 3-5 on the same N ranks, after the headline run: FSDP on the ~1.3 B GPT, the GPipe /
    1F1B pipeline (stages = N up to 4) and pipeline x DDP (2 stages x N/2).  Same timing rules as the headline
    (tools/bench_recipes.py: warm-up, barrier + synchronize, CUDA events, max over ranks, AdamW inside)."""
    sys.path.insert(0, str(ROOT / "tools"))
    from bench_recipes import run_recipe
    plan = [("fsdp_1p3b", dict(recipe="fsdp", model="1.3b", steps=8, warmup=5, measure_exposed=impl == "ours"))]
    if n <= 4:
        plan.append((f"pipe_{n}stage_m{n}", dict(recipe="pipe", steps=20, warmup=6, schedule="1f1b", microbatches=n)))
        plan.append((f"pipe_{n}stage_m{2 * n}", dict(recipe="pipe", steps=20, warmup=6, schedule="1f1b", microbatches=2 * n)))
    if n >= 4:
        plan.append((f"pipe_ddp_2x{n // 2}", dict(recipe="pipe-ddp", steps=20, warmup=6, schedule="1f1b", microbatches=4)))
    out = {}
    t_start = time.time()
    for name, kw in plan:
        if args.recipes != "all" and name.split("_")[0] not in args.recipes.split(","):
            continue
        # the recipes must never cost the headline line: stop adding them once the time budget is used up (every rank
        # takes the same decision: rank 0's clock is broadcast)
        spent = torch.tensor([time.time() - t_start], device=f"cuda:{torch.cuda.current_device()}")
        dist.broadcast(spent, src=0)
        if float(spent) > args.recipe_budget_s:
            out[name] = {"skipped": f"recipe time budget of {args.recipe_budget_s:.0f} s used up"}
            continue
        try:
            r = run_recipe(impl=impl, **kw)
            out[name] = {k: r.get(k) for k in ("ms_per_step", "tokens_per_s", "loss_after", "comm", "stages", "dp",
                                               "microbatches", "schedule", "global_batch", "seq_len", "model",
                                               "compute_only_ms_per_step", "exposed_comm_ms_per_step", "exposed_comm_error")
                         if r.get(k) is not None}
        except Exception as e:  # noqa: BLE001
            out[name] = {"error": repr(e)[:300]}
            try:
                dist.barrier()
            except Exception:
                pass
    return out


# ------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's own model + training step (torch autocast + GradScaler + AdamW, DDP for
    N > 1, torch.compile as in its default), from the verbatim copy under baseline/_ref."""
    from baseline.harness import load_reference, reference_unavailable_reason
    why = reference_unavailable_reason()
    rank, local_rank, n = world()
    if why is not None:
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)
        return
    import torch.nn.functional as F
    device = torch.device(f"cuda:{local_rank}")
    torch.cuda.set_device(device)
    if n > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=device)
    ref = load_reference()
    batches = host_batches(max(4, min(args.steps, 16)), seed=100 + rank)
    pad = 2
    results = {}
    modes = {"both": [True, False], "on": [True], "off": [False]}[args.ref_compile]
    clocks = None
    for use_compile in modes:
        torch.manual_seed(0)
        sp_model = ref.gpt.TransformerDecoderLM(dim=CFG["dim"], head_dim=CFG["head_dim"], heads=CFG["heads"],
                                                num_layers=CFG["num_layers"], vocab_size=CFG["vocab_size"],
                                                max_position_embeddings=CFG["sequence_length"]).to(device)
        model = sp_model
        if n > 1:
            from torch.nn.parallel import DistributedDataParallel as DDP
            model = DDP(sp_model, device_ids=[device.index])
        model.train()
        compiled_ok = False
        if use_compile:
            try:
                model = torch.compile(model)
                compiled_ok = True
            except Exception:
                compiled_ok = False
        optim = torch.optim.AdamW(model.parameters(), lr=CFG["learning_rate"])
        scaler = torch.amp.GradScaler("cuda", enabled=True)

        def step(batch, targets, read_loss):
            optim.zero_grad()
            with torch.autocast(device_type="cuda", dtype=torch.bfloat16, enabled=True):
                logits = model(**batch)
                logits, targets = logits.view(-1, sp_model.vocab_size), targets.view(-1)
                loss = F.cross_entropy(logits, targets, ignore_index=-100)
            scaler.scale(loss).backward()
            scaler.step(optim)
            scaler.update()
            return loss.item() if read_loss else loss

        try:
            for i in range(max(args.warmup, 5)):
                b, t = ref.utils.prepare_batch(batches[i % len(batches)], pad, device)
                step(b, t, True)
        except Exception as e:  # e.g. inductor unavailable on the box
            if use_compile:
                results[f"compile_error"] = repr(e)[:200]
                continue
            raise
        torch.cuda.synchronize(device)
        dev_batches = [ref.utils.prepare_batch(b, pad, device) for b in batches]
        sampler = ClockSampler(local_rank) if rank == 0 else None
        dev_blocks = timed_blocks(lambda i: step(*dev_batches[i % len(dev_batches)], False), args.steps, device,
                                  sampler, args.min_seconds)
        ms_dev = statistics.median(dev_blocks)
        if sampler is not None:
            clocks = sampler.summary()

        def e2e(i):
            b, t = ref.utils.prepare_batch(batches[i % len(batches)], pad, device)
            step(b, t, True)
        ms_e2e = statistics.median(timed_blocks(e2e, args.steps, device, None, args.min_seconds))
        results["compile" if use_compile and compiled_ok else "eager"] = (ms_dev, ms_e2e, len(dev_blocks))
        del model, optim, sp_model
        torch.cuda.empty_cache()

    tokens_per_step = CFG["batch_size"] * (CFG["sequence_length"] - 1) * n
    timed_modes = {k: v for k, v in results.items() if isinstance(v, tuple)}
    best = min(timed_modes, key=lambda k: timed_modes[k][0])
    ms_dev, ms_e2e, nblocks = timed_modes[best]
    extra = {"reference_mode": best,
             "all_modes_ms_per_step": {k: v[0] / args.steps for k, v in timed_modes.items()},
             "repairs_applied_in_memory": ref.repairs, "compile_error": results.get("compile_error"),
             "timed_blocks": nblocks, "e2e_ms_per_step": ms_e2e / args.steps}
    h2d = 3 * CFG["batch_size"] * (CFG["sequence_length"] - 1) * 8 + CFG["batch_size"] * (CFG["sequence_length"] - 1)
    line = result_line("reference", tokens_per_step * args.steps / ms_dev * 1e3, ms_dev / args.steps,
                       tokens_per_step * args.steps / ms_e2e * 1e3, n, args.steps, args.warmup, clocks,
                       0, extra, h2d, 4)
    if n > 1 and not args.no_recipes:
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        line["recipes"] = run_recipes("reference", n, args)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--ref-compile", dest="ref_compile", choices=["both", "on", "off"], default="both")
    ap.add_argument("--bucket_mb", type=float, default=16.0)
    ap.add_argument("--comm", choices=["auto", "symm", "nccl"], default="auto")
    ap.add_argument("--comm_dtype", choices=["fp32", "bf16"], default="fp32")
    ap.add_argument("--min-seconds", dest="min_seconds", type=float, default=1.0,
                    help="repeat the K-step timed block until this much device time has been measured (median reported)")
    ap.add_argument("--no-recipes", dest="no_recipes", action="store_true",
                    help="N > 1: skip the FSDP-1.3B / pipeline / pipeline x DDP recipes after the headline run")
    ap.add_argument("--recipes", default="all", help="comma list out of fsdp,pipe (pipe covers pipe and pipe_ddp)")
    ap.add_argument("--no-selfcheck", dest="no_selfcheck", action="store_true")
    ap.add_argument("--recipe-budget-s", dest="recipe_budget_s", type=float, default=240.0,
                    help="wall-clock budget for the recipes of one arm; recipes that would start later are skipped")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    rank, _, n = world()
    if args.gpus > 1 and n == 1:
        # convenience: re-launch ourselves under torchrun when started as a plain process
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 400), __file__,
               *sys.argv[1:]]
        raise SystemExit(subprocess.call(cmd))
    if not torch.cuda.is_available():
        print(json.dumps({"impl": args.impl, "unavailable": "no CUDA device on this box"}))
        return
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
