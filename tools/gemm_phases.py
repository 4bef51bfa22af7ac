"""Per-CTA phase timestamps and wait accounting of the tcgen05 GEMM (bring-up): where does the
time of a GEMM go - operands late (TMA / L2), accumulator not drained (epilogue), or MMA?

    python tools/gemm_phases.py            # all shapes, single-CTA and CTA-pair variants
    ONLY_LM=1 python tools/gemm_phases.py  # LM-head shapes only
"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402

from distributed_pytorch_cookbook_b200.ops import _lib, native  # noqa: E402

BF16 = torch.bfloat16
names = ["entry", "setup_done", "first_tma_issued", "first_kblock_landed", "tile0_mma_committed",
         "tile0_acc_ready(epi)", "epi_last_store_issued", "epi_stores_done", "pre_dealloc", "dealloc_done", "producer_done"]
CASES = [("lm fwd 4096", 4096, 50257, 256, False, 0, 0, False), ("wgrad lm 4096", 50257, 256, 4096, False, 1, 1, True),
         ("dgrad lm 4096", 4096, 256, 50257, False, 0, 1, True), ("fwd qkv", 16320, 768, 256, False, 0, 0, False),
         ("fwd up", 16320, 1024, 256, False, 0, 0, False), ("dgrad qkv", 16320, 256, 768, False, 0, 1, False),
         ("wgrad up", 1024, 256, 16320, False, 1, 1, True), ("wgrad qkv", 768, 256, 16320, False, 1, 1, True),
         ("dgrad down(mask)", 16320, 1024, 256, False, 0, 1, False)]
if os.environ.get("ONLY_LM"):
    CASES = CASES[:3]
VARIANTS = [("1sm", 0x800), ("2sm", 0x400)]


def run(name, M, N, K, res, a_mn, b_mn, acc, vbit):
    Kp = (K + 7) // 8 * 8
    A = torch.randn(K, (M + 7) // 8 * 8, device="cuda").to(BF16) if a_mn else torch.randn(M, Kp, device="cuda").to(BF16)
    W = torch.randn(K, (N + 7) // 8 * 8, device="cuda").to(BF16) if b_mn else torch.randn(N, Kp, device="cuda").to(BF16)
    ld = (N + 127) // 128 * 128
    out = torch.zeros(M, ld, device="cuda", dtype=torch.float32 if (res or acc) else BF16)
    r = torch.randn(M, ld, device="cuda") if res else None
    dbg = torch.zeros(4096 * 32, dtype=torch.int64, device="cuda")
    kw = dict(residual=r, a_mn_major=bool(a_mn), b_mn_major=bool(b_mn), accumulate=acc, cfg=vbit)
    for _ in range(3):
        native.gemm(A, W, out, M, N, K, **kw)
    torch.cuda.synchronize()
    _lib.call("b200_gemm_set_debug_buffer", dbg.data_ptr())
    native.gemm(A, W, out, M, N, K, **kw)
    torch.cuda.synchronize()
    _lib.call("b200_gemm_set_debug_buffer", None)
    d = dbg.view(-1, 32).cpu()
    used = d[d[:, 0] > 0]
    print(f"\n{name}: M={M} N={N} K={K}: CTAs={len(used)}")
    rel = (used - used[:, :1]).float()
    for i, n in enumerate(names):
        col = rel[:, i]
        print(f"  {n:26s} median {col.median().item():9.0f} cyc   max {col.max().item():9.0f}")
    life = (used[:, 9] - used[:, 0]).float().median().item()
    lead = used[(used[:, 14] + used[:, 15]) > 0]          # CTAs that issued MMAs (the leaders of a pair)
    if len(lead):
        print(f"  wait accounting, median cycles per CTA (CTA life {life:.0f}): "
              f"MMA<-operands {lead[:, 14].float().median():.0f} | MMA<-accumulator drained {lead[:, 15].float().median():.0f} | "
              f"epilogue<-MMA {used[:, 16].float().median():.0f} | epilogue<-staging buffer {used[:, 18].float().median():.0f} | "
              f"producer<-free slot {used[:, 17].float().median():.0f}")
    t0 = used[:, 12].min()
    ent = (used[:, 12] - t0).float()
    ext = (used[:, 13] - t0).float()
    print(f"  globaltimer: entry min/median/max {ent.min():.0f}/{ent.median():.0f}/{ent.max():.0f} ns; "
          f"exit median/max {ext.median():.0f}/{ext.max():.0f} ns; CTA life median {(ext - ent).median():.0f} ns")


for (name0, M, N, K, res, a_mn, b_mn, acc) in CASES:
    for vname, vbit in VARIANTS:
        run(f"{name0} [{vname}]", M, N, K, res, a_mn, b_mn, acc, vbit)
