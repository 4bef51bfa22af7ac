"""Where the cycles of the warp-specialised attention backward go (trace build: B200_GEMM_TRACE=1 python -m csrc.build):
per CTA, the MMA warp's time in each wait and the first compute warp's time in its waits / math / dQ drain."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from distributed_pytorch_cookbook_b200.ops import _lib
B, T, H, D = 8, 1023, 16, 128
if len(sys.argv) > 1:
    B, T, H, D = (int(x) for x in sys.argv[1:5])
mode = int(sys.argv[5]) if len(sys.argv) > 5 else 1
qkv = (torch.randn(B, T, 3 * H * D, device="cuda") * 0.7).to(torch.bfloat16)
out = torch.empty((B, T, H * D), dtype=torch.bfloat16, device="cuda")
lse = torch.empty((B, H, T), dtype=torch.float32, device="cuda")
dout = torch.randn_like(out); dqkv = torch.empty_like(qkv)
delta = torch.empty((B, H, T), dtype=torch.float32, device="cuda")
dq_acc = torch.empty((B, T, H * D), dtype=torch.float32, device="cuda")
sc = D ** -0.5
_lib.call("b200_attn_fwd", qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), B, T, H, D, sc, None, _lib.stream())
def bwd():
    _lib.call("b200_attn_bwd_select", qkv.data_ptr(), out.data_ptr(), dout.data_ptr(), lse.data_ptr(), dqkv.data_ptr(),
              delta.data_ptr(), dq_acc.data_ptr(), B, T, H, D, sc, None, mode, _lib.stream())
for _ in range(3):
    bwd()
torch.cuda.synchronize()
nkt = (T + 127) // 128
dbg = torch.zeros(nkt * H * B * 32, dtype=torch.int64, device="cuda")
_lib.call("b200_attn_set_debug_buffer", dbg.data_ptr())
bwd()
torch.cuda.synchronize()
_lib.call("b200_attn_set_debug_buffer", None)
d = dbg.view(B, H, nkt, 32).float()          # grid (key tiles, H, B)
names = {0: "MMA: wait Q/dO loaded", 1: "MMA: wait dQ drained", 2: "MMA: wait P^T/dS^T ready", 3: "MMA: warp lifetime",
         8: "compute: wait stats staged", 9: "compute: wait S^T/dP^T", 10: "compute: wait dQ MMA", 11: "compute: math + stores",
         12: "compute: dQ drain"}
print(f"B{B} T{T} H{H} D{D} mode {mode}: cycles per (query tile, key tile) pair, mean over CTAs")
its = d[..., 4].clamp(min=1)
for k, n in names.items():
    print(f"  {n:30s} {(d[..., k] / its).mean().item():9.0f}")
