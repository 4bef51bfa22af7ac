"""Per-CTA phase timestamps of the tcgen05 GEMM (bring-up): where does a tile's time go?"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from distributed_pytorch_cookbook_b200.ops import _lib, native

BF16 = torch.bfloat16
names = ["entry", "setup_done", "first_tma_issued", "first_kblock_landed", "tile0_mma_committed",
         "tile0_acc_ready(epi)", "epi_last_store_issued", "epi_stores_done", "pre_dealloc", "dealloc_done", "producer_done"]
for (M, N, K, res) in [(16320, 768, 256, False), (16320, 256, 256, True), (16320, 1024, 256, False), (1024, 50257, 256, False)]:
    A = torch.randn(M, K, device="cuda").to(BF16); W = torch.randn(N, K, device="cuda").to(BF16)
    ld = (N + 127) // 128 * 128
    out = torch.empty(M, ld, device="cuda", dtype=torch.float32 if res else BF16)
    r = torch.randn(M, ld, device="cuda") if res else None
    dbg = torch.zeros(4096 * 16, dtype=torch.int64, device="cuda")
    for it in range(3):
        native.gemm(A, W, out, M, N, K, residual=r)
    torch.cuda.synchronize()
    _lib.call("b200_gemm_set_debug_buffer", dbg.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); native.gemm(A, W, out, M, N, K, residual=r); e1.record()
    torch.cuda.synchronize()
    _lib.call("b200_gemm_set_debug_buffer", None)
    d = dbg.view(-1, 16).cpu()
    used = d[d[:, 0] > 0]
    print(f"\nM={M} N={N} K={K} residual={res}: {e0.elapsed_time(e1)*1e3:.1f} us, CTAs={len(used)}")
    rel = (used - used[:, :1]).float()
    for i, n in enumerate(names):
        col = rel[:, i]
        print(f"  {n:26s} median {col.median().item():9.0f} cyc   max {col.max().item():9.0f}")
    span = (used[:, 9].max() - used[:, 0].min()).item()
    print(f"  whole-kernel span (first entry -> last dealloc): {span} cycles; entry spread {int((used[:,0].max()-used[:,0].min()).item())}")
