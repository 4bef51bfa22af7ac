"""Benchmarks for the BASELINE.md rows that bench.py (row 2, DDP) does not cover, launched under
torchrun:  --recipe fsdp | pipe | pipe-ddp   --impl ours | reference   [--model default|1.3b]

Same timing discipline as bench.py: warm-up, barrier + synchronize on both sides, CUDA events, max
over ranks, AdamW step inside the timed region, synthetic tokens, random init.  The reference arm
runs the reference's own model from baseline/_ref through torch FSDP (main-fsdp.py:60-69 with the
kwarg repair R4) or torch.distributed.pipelining (the R6 re-expression of main-pipe.py's intent).
Prints one JSON line on rank 0.
"""
import argparse
import functools
import json
import os
import sys
from pathlib import Path
from types import SimpleNamespace

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import torch.nn.functional as F  # noqa: E402

MODELS = {
    "default": dict(dim=256, head_dim=32, heads=8, num_layers=8, sequence_length=256, vocab_size=50257, batch_size=64),
    "1.3b": dict(dim=2048, head_dim=128, heads=16, num_layers=24, sequence_length=1024, vocab_size=50257, batch_size=8),
}


def timed(fn, steps, dev):
    dist.barrier(); torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record(); torch.cuda.synchronize(dev); dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms)


def run_recipe(recipe, impl="ours", model="default", steps=10, warmup=4, schedule="1f1b", microbatches=0,
               comm="auto", no_graph=False, profile_steps=False, measure_exposed=False):
    """One recipe on the already-initialised default process group (device = cuda:LOCAL_RANK).
    Returns the result dict on every rank (timing is the max over ranks)."""
    import gc
    rank, world = dist.get_rank(), dist.get_world_size()
    dev = torch.device(f"cuda:{torch.cuda.current_device()}")
    a = SimpleNamespace(recipe=recipe, impl=impl, model=model, steps=steps, warmup=warmup, schedule=schedule,
                        microbatches=microbatches, comm=comm, no_graph=no_graph, profile=profile_steps)
    cfg = MODELS[a.model]
    B, S, V = cfg["batch_size"], cfg["sequence_length"], cfg["vocab_size"]
    dp = world if a.recipe == "fsdp" else (world // 2 if a.recipe == "pipe-ddp" else 1)
    stages = 1 if a.recipe == "fsdp" else world // dp
    M = a.microbatches or stages
    g = torch.Generator().manual_seed(10 + (rank % dp if a.recipe != "pipe" else 0))
    ids = [torch.randint(0, V, (B, S - 1), generator=g).to(dev) for _ in range(4)]
    pos = torch.arange(S - 1, device=dev).unsqueeze(0).repeat(B, 1)
    extra = {}
    st = None
    torch.manual_seed(0)
    if a.impl == "ours":
        from distributed_pytorch_cookbook_b200.trainer import build_model
        ns = SimpleNamespace(**cfg, learning_rate=1e-4, disable_amp=False, disable_cuda_graph=a.no_graph, bucket_mb=16.0,
                             comm=a.comm, cpu_offload=False)
        model = build_model(ns, V)
        if a.recipe == "fsdp":
            from distributed_pytorch_cookbook_b200.parallel.fsdp import FSDPStrategy
            st = FSDPStrategy()
        else:
            from distributed_pytorch_cookbook_b200.parallel.pipe import PipeStrategy
            st = PipeStrategy(num_stages=stages, num_microbatches=M, schedule=a.schedule, dp_size=dp)
        st.setup(model, ns, dev)
        st.train()
        step = lambda i: st.train_step({"input_ids": ids[i % 4], "position_ids": pos}, ids[i % 4], amp=True)  # noqa: E731
        extra["comm"] = st.comm_kind
    else:
        from baseline.harness import load_reference
        ref = load_reference()
        sp = ref.gpt.TransformerDecoderLM(dim=cfg["dim"], head_dim=cfg["head_dim"], heads=cfg["heads"],
                                          num_layers=cfg["num_layers"], vocab_size=V, max_position_embeddings=S)
        if a.recipe == "fsdp":
            from torch.distributed.fsdp import FullyShardedDataParallel as FSDP
            from torch.distributed.fsdp.wrap import size_based_auto_wrap_policy
            model = FSDP(sp.to(dev), device_id=dev,
                         auto_wrap_policy=functools.partial(size_based_auto_wrap_policy, min_num_params=100))
            optim = torch.optim.AdamW(model.parameters(), lr=1e-4)
            scaler = torch.amp.GradScaler("cuda")

            def step(i):
                optim.zero_grad()
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    logits = model(input_ids=ids[i % 4], position_ids=pos)
                    loss = F.cross_entropy(logits.view(-1, V), ids[i % 4].view(-1), ignore_index=-100)
                scaler.scale(loss).backward()
                scaler.step(optim); scaler.update()
                return loss
        else:
            # R6: the reference's intent (main-pipe.py:52-83) on torch.distributed.pipelining
            from torch.distributed.pipelining import PipelineStage, Schedule1F1B, ScheduleGPipe
            sys.path.insert(0, str(ROOT))
            from distributed_pytorch_cookbook_b200.parallel.pipe import stage_layer_range
            pp_rank, dp_rank = rank // dp, rank % dp
            lo, hi = stage_layer_range(cfg["num_layers"], stages, pp_rank)

            class Stage(torch.nn.Module):
                def __init__(self):
                    super().__init__()
                    self.emb = sp.embeddings if pp_rank == 0 else None
                    self.layers = torch.nn.ModuleList([sp.decoder.layers[i] for i in range(lo, hi)])
                    self.norm = sp.norm_out if pp_rank == stages - 1 else None
                    self.head = sp.lm_head if pp_rank == stages - 1 else None

                def forward(self, x):
                    with torch.autocast("cuda", dtype=torch.bfloat16):
                        if self.emb is not None:
                            x = self.emb(x, pos[: x.shape[0]])
                        for l in self.layers:
                            x = l(x)
                        if self.head is not None:
                            x = self.head(self.norm(x))
                    return x
            mod = Stage().to(dev)
            pp_group = dist.new_group([s * dp + dp_rank for s in range(stages)]) if dp > 1 else None
            if dp > 1:
                for s in range(stages):
                    for d in range(dp):
                        if d != dp_rank:
                            dist.new_group([t * dp + d for t in range(stages)])
                dgroups = [dist.new_group([s * dp + d for d in range(dp)]) for s in range(stages)]
                from torch.nn.parallel import DistributedDataParallel as DDP
                mod = DDP(mod, device_ids=[dev.index], process_group=dgroups[pp_rank])
            stage = PipelineStage(mod, pp_rank, stages, dev, group=pp_group)
            loss_fn = lambda out, tgt: F.cross_entropy(out.reshape(-1, V).float(), tgt.reshape(-1))  # noqa: E731
            sched = (ScheduleGPipe if a.schedule == "gpipe" else Schedule1F1B)(stage, M, loss_fn=loss_fn)
            optim = torch.optim.AdamW(mod.parameters(), lr=1e-4)

            def step(i):
                optim.zero_grad()
                ls = []
                if pp_rank == 0:
                    sched.step(ids[i % 4])
                elif pp_rank == stages - 1:
                    sched.step(target=ids[i % 4], losses=ls)
                else:
                    sched.step()
                optim.step()
                return torch.stack([l.detach().float() for l in ls]).mean() if ls else torch.full((), -1e30, device=dev)
    last = None
    for i in range(a.warmup):
        last = step(i)
    if a.profile:
        from torch.profiler import ProfilerActivity, profile
        torch.cuda.synchronize(dev); dist.barrier()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for i in range(3):
                step(i)
            torch.cuda.synchronize(dev)
        if rank == 0:
            if os.environ.get("TRACE_OUT"):     # chrome trace of the 3 profiled steps (per-stream timeline)
                prof.export_chrome_trace(os.environ["TRACE_OUT"])
            rows = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)
            tot = sum(e.device_time_total for e in rows)
            print(f"total device time {tot/3:.1f} us/step (sum over streams)")
            for e in rows[:16]:
                print(f"{e.device_time_total/3:10.1f} us/step n/step={e.count/3:6.1f} avg={e.device_time_total/e.count:9.1f}  {e.key[:100]}")
        dist.barrier()
    ms = timed(step, a.steps, dev)
    last = step(0)
    loss = None
    if torch.is_tensor(last):
        lt = last.detach().float().reshape(1).clone()
        dist.all_reduce(lt, op=dist.ReduceOp.MAX)       # pipeline: only some ranks hold it; NaN propagates
        loss = float(lt)
    tokens = B * (S - 1) * dp
    out = {"recipe": a.recipe, "impl": a.impl, "model": a.model, "n_gpus": world, "stages": stages, "dp": dp,
           "microbatches": M, "schedule": a.schedule if a.recipe != "fsdp" else None,
           "ms_per_step": ms / a.steps, "tokens_per_s": tokens * a.steps / ms * 1e3,
           "global_batch": B * dp, "seq_len": S, "loss_after": loss, **extra}
    if measure_exposed and st is not None and hasattr(st, "set_skip_comm"):
        # the same step with every collective skipped (each rank updates its shard from its local gradients):
        # what is left of the difference is communication that compute did not hide
        try:
            st.set_skip_comm(True)
            for i in range(6):                       # eager warm-up steps + the (re-)capture of the step graph
                step(i)
            ms_nc = timed(step, a.steps, dev)
            out["compute_only_ms_per_step"] = ms_nc / a.steps
            out["exposed_comm_ms_per_step"] = max(ms - ms_nc, 0.0) / a.steps
        except Exception as e:  # noqa: BLE001
            out["exposed_comm_error"] = repr(e)[:200]
    if st is not None and hasattr(st, "close"):
        st.close()
    del st, step
    gc.collect()
    torch.cuda.empty_cache()
    dist.barrier()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--recipe", choices=["fsdp", "pipe", "pipe-ddp"], required=True)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--model", choices=list(MODELS), default="default")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--schedule", default="1f1b")
    ap.add_argument("--microbatches", type=int, default=0)
    ap.add_argument("--comm", default="auto")
    ap.add_argument("--no_graph", action="store_true")
    ap.add_argument("--profile", action="store_true")
    ap.add_argument("--exposed", action="store_true")
    a = ap.parse_args()
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device(f"cuda:{local}")
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    out = run_recipe(a.recipe, a.impl, a.model, a.steps, a.warmup, a.schedule, a.microbatches, a.comm, a.no_graph,
                     a.profile, a.exposed)
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
