"""The kernels of the 1.4 B configuration (and the fused LM head of the default one), each launched once or twice,
as the command ncu wraps (one GPU):  attention forward / backward at d = 128, T = 1023, B = 8, H = 16; the layer GEMMs
of dim 2048 (QKV, out-projection + residual, MLP up + ReLU, MLP down, masked dgrad + bias gradient, wgrad); the
LM-head softmax-epilogue GEMM, its finish kernel and the row-scaled dX GEMM at 16320 x 50257 x 256.

    ncu --set full --clock-control none --import-source on -o gpurun_out/targets python tools/ncu_targets.py
"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402

from distributed_pytorch_cookbook_b200.ops import _lib, native  # noqa: E402

BF = torch.bfloat16
dev = torch.device("cuda")
torch.manual_seed(0)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2

# ---- attention, 1.4 B shape -------------------------------------------------------------------------------
B, T, H, D = 8, 1023, 16, 128
qkv = (torch.randn(B, T, 3 * H * D, device=dev) * 0.7).to(BF).requires_grad_(True)
for _ in range(reps):
    qkv.grad = None
    out = native.attention(qkv, H, D, D ** -0.5, None)
    out.backward(torch.randn_like(out))

# ---- layer GEMMs, dim 2048, 8184 tokens ---------------------------------------------------------------------
M, dim = 8184, 2048
x = torch.randn(M, dim, device=dev).to(BF).requires_grad_(True)
res = torch.randn(M, dim, device=dev).requires_grad_(True)
wq, wk, wv = (torch.nn.Parameter(torch.randn(dim, dim, device=dev) * 0.02) for _ in range(3))
wo = torch.nn.Parameter(torch.randn(dim, dim, device=dev) * 0.02)
bo = torch.nn.Parameter(torch.zeros(dim, device=dev))
w1 = torch.nn.Parameter(torch.randn(4 * dim, dim, device=dev) * 0.02)
b1 = torch.nn.Parameter(torch.zeros(4 * dim, device=dev))
w2 = torch.nn.Parameter(torch.randn(dim, 4 * dim, device=dev) * 0.01)
b2 = torch.nn.Parameter(torch.zeros(dim, device=dev))
for _ in range(reps):
    y = native.linear_qkv(x, wq, wk, wv)
    y.backward(torch.randn_like(y))
    o = native.linear(x, wo, bo, residual=res, out_fp32=True)
    o.backward(torch.randn_like(o))
    m = native.mlp(x, w1, b1, w2, b2, residual=res)
    m.backward(torch.randn_like(m))

# ---- LM head, default configuration ---------------------------------------------------------------------------
Ml, Dl, V = 16320, 256, 50257
xl = torch.randn(Ml, Dl, device=dev).to(BF).requires_grad_(True)
wl = torch.nn.Parameter(torch.randn(V, Dl, device=dev) * 0.036)
tgt = torch.randint(0, V, (Ml,), device=dev)
for _ in range(reps):
    loss, _, _ = native.lmhead_cross_entropy(xl, wl, tgt)
    loss.backward()
torch.cuda.synchronize()
print("done", float(loss), _lib.launch_count())
