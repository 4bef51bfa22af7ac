"""Per-CTA phase timestamps of the attention backward kernel (bring-up; needs the trace build: B200_GEMM_TRACE=1 python -m csrc.build): default GPT shape."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from distributed_pytorch_cookbook_b200.ops import _lib, native
B, T, H, D = 64, 255, 8, 32
qkv = (torch.randn(B, T, 3 * H * D, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_(True)
dout = torch.randn(B, T, H * D, device="cuda").to(torch.bfloat16)
for _ in range(3):
    native.attention(qkv, H, D, D ** -0.5).backward(dout)
torch.cuda.synchronize()
nkt = (T + 127) // 128
dbg = torch.zeros(nkt * H * B * 32, dtype=torch.int64, device="cuda")
out = native.attention(qkv, H, D, D ** -0.5)
torch.cuda.synchronize()
_lib.call("b200_attn_set_debug_buffer", dbg.data_ptr())
out.backward(dout)
torch.cuda.synchronize()
_lib.call("b200_attn_set_debug_buffer", None)
d = dbg.view(nkt, B, H, 32).cpu()      # grid (H, B, key tiles): index ((z * gy + y) * gx + x)
names = {0: "entry", 1: "setup + K/V loaded", 2: "it0 Q/dO loaded", 3: "it0 S ready", 4: "it0 P written", 5: "it0 dP ready",
         6: "it0 dS written", 7: "it0 dV/dK/dQ MMAs done", 8: "it0 dQ reduced", 10: "it1 Q/dO loaded", 11: "it1 S ready",
         12: "it1 P written", 13: "it1 dP ready", 14: "it1 dS written", 15: "it1 MMAs done", 16: "it1 dQ reduced",
         20: "loop end", 21: "dK/dV written + dealloc"}
for kt in range(nkt):
    rows = d[kt].reshape(-1, 32)
    rel = (rows - rows[:, :1]).float()
    print(f"key tile {kt} ({nkt - kt} query tiles):")
    for k, n in names.items():
        if rows[:, k].max() > 0:
            print(f"  {n:26s} median {rel[:, k].median().item():9.0f} cyc")
allr = d.reshape(-1, 32)
t0 = allr[:, 24].min()
ent, ext = (allr[:, 24] - t0).float(), (allr[:, 25] - t0).float()
print(f"globaltimer: entry median/max {ent.median():.0f}/{ent.max():.0f} ns, exit median/max {ext.median():.0f}/{ext.max():.0f} ns, "
      f"CTA life median {(ext - ent).median():.0f} max {(ext - ent).max():.0f} ns")
