"""Per-CTA phase timestamps of the tcgen05 GEMM (bring-up): where does a tile's time go?"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from distributed_pytorch_cookbook_b200.ops import _lib, native

BF16 = torch.bfloat16
names = ["entry", "setup_done", "first_tma_issued", "first_kblock_landed", "tile0_mma_committed",
         "tile0_acc_ready(epi)", "epi_last_store_issued", "epi_stores_done", "pre_dealloc", "dealloc_done", "producer_done"]
CASES = [("lm fwd 4096", 4096, 50257, 256, False, 0, 0, False), ("wgrad lm 4096", 50257, 256, 4096, False, 1, 1, True),
         ("dgrad lm 4096", 4096, 256, 50257, False, 0, 1, True), ("fwd qkv", 16320, 768, 256, False, 0, 0, False),
         ("fwd up", 16320, 1024, 256, False, 0, 0, False), ("dgrad qkv", 16320, 256, 768, False, 0, 1, False),
         ("wgrad up", 1024, 256, 16320, False, 1, 1, True), ("wgrad qkv", 768, 256, 16320, False, 1, 1, True),
         ("dgrad down(mask)", 16320, 1024, 256, False, 0, 1, False)]
for (name, M, N, K, res, a_mn, b_mn, acc) in CASES:
    Kp = (K + 7) // 8 * 8
    A = torch.randn(K, (M + 7) // 8 * 8, device="cuda").to(BF16) if a_mn else torch.randn(M, Kp, device="cuda").to(BF16)
    W = torch.randn(K, (N + 7) // 8 * 8, device="cuda").to(BF16) if b_mn else torch.randn(N, Kp, device="cuda").to(BF16)
    ld = (N + 127) // 128 * 128
    out = torch.zeros(M, ld, device="cuda", dtype=torch.float32 if (res or acc) else BF16)
    r = torch.randn(M, ld, device="cuda") if res else None
    dbg = torch.zeros(4096 * 32, dtype=torch.int64, device="cuda")
    kw = dict(residual=r, a_mn_major=bool(a_mn), b_mn_major=bool(b_mn), accumulate=acc, cfg=256 if "bn256" in name else 0)
    for it in range(3):
        native.gemm(A, W, out, M, N, K, **kw)
    torch.cuda.synchronize()
    _lib.call("b200_gemm_set_debug_buffer", dbg.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); native.gemm(A, W, out, M, N, K, **kw); e1.record()
    torch.cuda.synchronize()
    _lib.call("b200_gemm_set_debug_buffer", None)
    d = dbg.view(-1, 32).cpu()
    used = d[d[:, 0] > 0]
    print(f"\n{name}: M={M} N={N} K={K}: CTAs={len(used)}")
    rel = (used - used[:, :1]).float()
    for i, n in enumerate(names):
        col = rel[:, i]
        print(f"  {n:26s} median {col.median().item():9.0f} cyc   max {col.max().item():9.0f}")
    for c in range(4):
        cols = rel[:, 16 + 4 * c:20 + 4 * c].median(dim=0).values.tolist()
        print(f"  chunk {c}: ld_done {cols[0]:.0f}  math_done {cols[1]:.0f}  staged {cols[2]:.0f}  store_issued {cols[3]:.0f}")
    t0 = used[:, 12].min()
    ent = (used[:, 12] - t0).float(); ext = (used[:, 13] - t0).float()
    print(f"  globaltimer: entry min/median/max {ent.min():.0f}/{ent.median():.0f}/{ent.max():.0f} ns; exit median/max {ext.median():.0f}/{ext.max():.0f} ns; CTA life median {(ext-ent).median():.0f} ns")
    span = (used[:, 9].max() - used[:, 0].min()).item()
    print(f"  whole-kernel span (first entry -> last dealloc): {span} cycles; entry spread {int((used[:,0].max()-used[:,0].min()).item())}")
