"""Race stress of the cross-GPU flag protocols (the stand-in for compute-sanitizer's racecheck, which cannot see
across GPUs anyway): many replays of the collective kernels while every rank delays itself by a different, random
amount before each launch, bucket sizes / offsets / channel ranges alternate, and every result is compared with
NCCL.  Then one forced failure: a rank skips a launch and the others must report a timeout through the device
error word (and the Trainer's check must raise) instead of hanging or returning silently.

    torchrun --nproc-per-node N tools/race_stress.py [iterations]

One JSON line per check on rank 0; exit code != 0 on any failure.  Results: profiles/race_stress_<N>gpu.jsonl.
"""
import json
import os
import random
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from distributed_pytorch_cookbook_b200.ops import _lib  # noqa: E402
from distributed_pytorch_cookbook_b200.runtime.symm import (CH_ALLGATHER, CH_ALLREDUCE, CH_RS_ADAMW,  # noqa: E402
                                                            SymmContext)

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device(f"cuda:{local}")
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 300
fails = []


def report(name, ok, **kw):
    if rank == 0:
        print(json.dumps({"check": name, "ok": bool(ok), "world": world, **kw}), flush=True)
    if not ok:
        fails.append(name)


def dev_err(clear=1):
    import ctypes
    v = ctypes.c_int(0)
    _lib.call("b200_device_error", ctypes.addressof(v), clear)
    return v.value


ctx = SymmContext(dev)
N = 4 * 1024 * 1024
grads = ctx.alloc(N, torch.float32)
stage = ctx.alloc(N, torch.bfloat16)
shard_n = N // world // 8 * 8
shards = ctx.alloc(shard_n, torch.bfloat16)
rng = random.Random(1234 + rank)          # per-rank delays
shape_rng = random.Random(99)             # same shapes on every rank
st = _lib.stream
bad = {"allreduce": 0, "allreduce_bf16": 0, "reduce_scatter": 0, "allgather": 0}
worst = dict.fromkeys(bad, 0.0)
full = torch.empty(shard_n * world, dtype=torch.bfloat16, device=dev)
rs_out = torch.empty(N, dtype=torch.float32, device=dev)
for it in range(iters):
    cnt = shape_rng.choice([4096, 65536, 1 << 20, N // 2]) // (8 * world) * (8 * world)
    off = shape_rng.randrange(0, (N - cnt) // 8 + 1) * 8
    nblocks = shape_rng.choice([1, 4, 16, 32])
    torch.manual_seed(it * 64 + rank)
    x = torch.randn(N, device=dev)
    ref = x[off:off + cnt].clone()
    dist.all_reduce(ref)
    ref /= world
    # --- fp32 all-reduce --------------------------------------------------------------------------------
    grads.tensor.copy_(x)
    _lib.call("b200_spin", rng.randrange(0, 2_000_000), st())
    _lib.call("b200_allreduce", grads.ptrs_dev, grads.mc_ptr or None, ctx.flags.ptrs_dev, rank, world, off, cnt,
              1.0 / world, nblocks, CH_ALLREDUCE, st())
    e = (grads.tensor[off:off + cnt] - ref).abs().max().item()
    worst["allreduce"] = max(worst["allreduce"], e)
    bad["allreduce"] += int(not (e < 1e-5))
    # --- bf16-wire all-reduce -----------------------------------------------------------------------------
    if stage.mc_ptr:
        g = x.clone()
        _lib.call("b200_spin", rng.randrange(0, 2_000_000), st())
        _lib.call("b200_allreduce_bf16", g.data_ptr(), stage.ptrs_dev, stage.mc_ptr, ctx.flags.ptrs_dev, rank, world,
                  off, cnt, 1.0 / world, nblocks, CH_ALLREDUCE, st())
        e = (g[off:off + cnt] - ref).abs().max().item()
        untouched = torch.equal(g[:off], x[:off]) and torch.equal(g[off + cnt:], x[off + cnt:])
        worst["allreduce_bf16"] = max(worst["allreduce_bf16"], e)
        bad["allreduce_bf16"] += int(not (e < 0.06 and untouched))      # bf16 rounding of N(0,1) addends and of the mean
    # --- reduce-scatter (small grid) ------------------------------------------------------------------------
    grads.tensor.copy_(x)
    per = cnt // world
    _lib.call("b200_spin", rng.randrange(0, 2_000_000), st())
    _lib.call("b200_reduce_scatter", grads.ptrs_dev, grads.mc_ptr or None, ctx.flags.ptrs_dev, rank, world,
              off + rank * per, per, 1.0 / world, rs_out.data_ptr(), 1, min(nblocks, 16), CH_RS_ADAMW, st())
    e = (rs_out[:per] - ref[rank * per:(rank + 1) * per]).abs().max().item()
    worst["reduce_scatter"] = max(worst["reduce_scatter"], e)
    bad["reduce_scatter"] += int(not (e < 1e-5))
    # --- all-gather ---------------------------------------------------------------------------------------
    shards.tensor.copy_(torch.randn(shard_n, device=dev))
    ref_full = torch.empty_like(full)
    dist.all_gather_into_tensor(ref_full, shards.tensor.clone())
    _lib.call("b200_spin", rng.randrange(0, 2_000_000), st())
    _lib.call("b200_allgather", shards.ptrs_dev, ctx.flags.ptrs_dev, rank, world, 0, shard_n * 2, full.data_ptr(), 1,
              min(nblocks, 32), CH_ALLGATHER, st())
    torch.cuda.synchronize()
    bad["allgather"] += int(not torch.equal(full, ref_full))
    dist.barrier()          # peers may still be reading this rank's buffers
err = dev_err()
for k in bad:
    t = torch.tensor([bad[k]], device=dev)
    dist.all_reduce(t)
    report(f"stress_{k}", int(t) == 0 and err == 0, iterations=iters, mismatches=int(t), max_err=worst[k], device_error=err,
           multicast=bool(grads.mc_ptr))

# --- forced timeout: the last rank does not launch; everyone else must see the error word, not a hang ---------------
_lib.call("b200_set_comm_timeout_ms", 300)
torch.cuda.synchronize(); dist.barrier()
if rank != world - 1:
    _lib.call("b200_allreduce", grads.ptrs_dev, grads.mc_ptr or None, ctx.flags.ptrs_dev, rank, world, 0, 4096,
              1.0 / world, 4, 200, st())
torch.cuda.synchronize()
code = dev_err(clear=0)
raised = False
try:
    from distributed_pytorch_cookbook_b200.trainer import check_device_error
    check_device_error("in the forced-timeout test")
except RuntimeError:
    raised = True
expect = rank != world - 1
t = torch.tensor([int((code in (301, 302)) == expect and raised == expect)], device=dev)
dist.all_reduce(t)
report("forced_timeout_is_reported", int(t) == world, code_rank0=code, trainer_raised_rank0=raised)
dev_err(clear=1)
dist.barrier()
dist.destroy_process_group()
sys.exit(1 if fails else 0)
