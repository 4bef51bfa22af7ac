"""Per-CTA phase timestamps of the attention forward kernel (bring-up; needs the trace build: B200_GEMM_TRACE=1 python -m csrc.build)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from distributed_pytorch_cookbook_b200.ops import _lib, native
B, T, H, D = 64, 255, 8, 32
qkv = (torch.randn(B, T, 3 * H * D, device="cuda") * 0.5).to(torch.bfloat16)
for _ in range(3):
    native.attention(qkv, H, D, D ** -0.5)
torch.cuda.synchronize()
nct = 2 * H * B
dbg = torch.zeros(nct * 32, dtype=torch.int64, device="cuda")
_lib.call("b200_attn_set_debug_buffer", dbg.data_ptr())
native.attention(qkv, H, D, D ** -0.5)
torch.cuda.synchronize()
_lib.call("b200_attn_set_debug_buffer", None)
d = dbg.view(-1, 32).cpu()
names = {0: "entry", 1: "setup+Q", 2: "it0 kv loaded", 3: "it0 S ready", 4: "it0 passA", 5: "it0 passB+sync", 6: "it0 PV ready",
         7: "it0 O upd", 8: "it1 kv loaded", 9: "it1 S ready", 10: "it1 passA", 11: "it1 passB+sync", 12: "it1 PV ready",
         13: "it1 O upd", 20: "end"}
for qt in (0, 1):
    rows = d[qt::2]
    rel = (rows - rows[:, :1]).float()
    print(f"q-tile {qt}:")
    for k, n in names.items():
        if rows[:, k].max() > 0:
            print(f"  {n:16s} median {rel[:, k].median().item():9.0f}")
t0 = d[:, 24].min()
ent, ext = (d[:, 24] - t0).float(), (d[:, 25] - t0).float()
print(f"globaltimer: entry median/max {ent.median():.0f}/{ent.max():.0f} ns, exit median/max {ext.median():.0f}/{ext.max():.0f} ns, life median {(ext-ent).median():.0f} max {(ext-ent).max():.0f}")
import numpy as np
e = np.sort(ent.numpy()); print("entry-time deciles (ns):", [int(e[int(i * (len(e) - 1) / 10)]) for i in range(11)])
