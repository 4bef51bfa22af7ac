import os, sys
from pathlib import Path
from types import SimpleNamespace
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch, torch.distributed as dist
from torch.profiler import ProfilerActivity, profile
rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device(f"cuda:{local}"); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
cfg = dict(dim=2048, head_dim=128, heads=16, num_layers=24, sequence_length=1024, vocab_size=50257, batch_size=8)
from distributed_pytorch_cookbook_b200.trainer import build_model
from distributed_pytorch_cookbook_b200.parallel.fsdp import FSDPStrategy
ns = SimpleNamespace(**cfg, learning_rate=1e-4, disable_amp=False, disable_cuda_graph=False, bucket_mb=4.0, comm="nccl", cpu_offload=False)
torch.manual_seed(0)
model = build_model(ns, 50257)
st = FSDPStrategy(); st.setup(model, ns, dev); st.train()
import time
from distributed_pytorch_cookbook_b200.ops import native
g = torch.Generator().manual_seed(10 + rank)
idl = [torch.randint(0, 50257, (8, 1023), generator=g).to(dev) for _ in range(4)]
pos = torch.arange(1023, device=dev).unsqueeze(0).repeat(8, 1)
cnt = [0]
def step():
    ids = idl[cnt[0] % 4]; cnt[0] += 1
    return st.train_step({"input_ids": ids, "position_ids": pos}, ids, amp=True)
for i in range(3):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    step()
    torch.cuda.synchronize(); t1 = time.perf_counter()
    if rank == 0: print(f"cpu step{i} wall {1e3*(t1-t0):.1f} ms lmstate {native.lm_state(model.lm_head.weight).tolist()}")
torch.cuda.synchronize(); dist.barrier()
for trial in range(14):
    ms0 = torch.cuda.memory_stats()
    t0 = time.perf_counter(); ts = []
    losses = []; flags = []
    for i in range(5):
        losses.append(step()); ts.append(time.perf_counter() - t0)

    torch.cuda.synchronize(); t1 = time.perf_counter()
    ms1 = torch.cuda.memory_stats()
    if rank == 0:
        ls = [float(l) for l in losses]
        print("cpu trial", trial, "NAN" if any(l != l for l in ls) else "ok", [round(l, 3) for l in ls])
        print(f"cpu nosync trial{trial}: total {1e3*(t1-t0)/5:.1f} ms/step, host enqueue times {[round(1e3*t,1) for t in ts]}, "
              f"segments +{ms1['segment.all.allocated']-ms0['segment.all.allocated']} retries +{ms1['num_alloc_retries']-ms0['num_alloc_retries']} "
              f"reserved {ms1['reserved_bytes.all.current']/2**30:.1f} GiB peak_alloc {ms1['allocated_bytes.all.peak']/2**30:.1f} GiB")
dist.barrier(); dist.destroy_process_group()
