"""Per-kernel device time of the graph-replayed flagship step via torch.profiler (CUPTI activity
tracing: kernels run back to back as in production, unlike ncu's serialised replays)."""
import os
import sys
from pathlib import Path
from types import SimpleNamespace

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402

from distributed_pytorch_cookbook_b200.trainer import SingleStrategy, build_model  # noqa: E402

BIG = os.environ.get("MODEL") == "1.3b"
B, S, V = (8, 1024, 50257) if BIG else (64, 256, 50257)
dev = torch.device("cuda")
torch.manual_seed(0)
ns = SimpleNamespace(dim=2048 if BIG else 256, head_dim=128 if BIG else 32, heads=16 if BIG else 8,
                     num_layers=24 if BIG else 8, sequence_length=S, learning_rate=1e-4,
                     disable_amp=False, disable_cuda_graph=os.environ.get("NOGRAPH") == "1")
model = build_model(ns, V)
st = SingleStrategy()
st.setup(model, ns, dev)
ids = torch.randint(0, V, (B, S - 1), device=dev)
pos = torch.arange(S - 1, device=dev).unsqueeze(0).repeat(B, 1)
batch = {"input_ids": ids, "position_ids": pos}
for _ in range(6):
    st.train_step(batch, ids)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(5):
        st.train_step(batch, ids)
    torch.cuda.synchronize()
rows = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)
tot = sum(e.device_time_total for e in rows)
print(f"total device time {tot/5:.1f} us/step")
for e in rows[:28]:
    print(f"{e.device_time_total/5:9.1f} us/step  {100*e.device_time_total/tot:5.1f}%  n/step={e.count/5:6.1f}  avg={e.device_time_total/e.count:7.1f}  {e.key[:90]}")
