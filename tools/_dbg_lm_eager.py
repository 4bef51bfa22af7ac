import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from distributed_pytorch_cookbook_b200.ops import native
torch.manual_seed(0)
M, D, V = 8184, 2048, 50257
x = torch.randn(M, D, device="cuda").to(torch.bfloat16).requires_grad_(True)
w = torch.nn.Parameter(torch.randn(V, D, device="cuda") * 0.0221 / 1.73 * 1.73)
tgt = torch.randint(0, V, (M,), device="cuda")
for it in range(6):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    loss, nv, nc = native.lmhead_cross_entropy(x, w, tgt)
    t1 = time.perf_counter()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    loss.backward()
    torch.cuda.synchronize(); t3 = time.perf_counter()
    print(f"it{it}: host fwd {1e3*(t1-t0):.2f} ms, fwd total {1e3*(t2-t0):.2f} ms, bwd {1e3*(t3-t2):.2f} ms  loss {float(loss):.4f} state {native.lm_state(w).tolist()}")
