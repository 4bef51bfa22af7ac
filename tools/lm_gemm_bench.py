"""Time the LM-head GEMMs on their own (CUDA events, 1 GPU).

    python tools/lm_gemm_bench.py [--rows 16320] [--dim 256] [--vocab 50257]

Prints one line per variant: the plain logits GEMM (bf16 out), the softmax-epilogue GEMM that stores
E = exp(logit - shift) and the statistics-only one (evaluation).  The output (rows x vocab bf16, 1.6 GB at
the default shape) is far larger than L2, so every iteration writes through to HBM.  Run once more with
B200GPT_GEMM_STAGING2=0 for the kernel with one staging set per epilogue warp.
"""
import argparse
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_pytorch_cookbook_b200.ops.native import call, gemm, ptr, stream  # noqa: E402


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) * 1e3 for i in range(iters))
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=16320)
    ap.add_argument("--dim", type=int, default=256)
    ap.add_argument("--vocab", type=int, default=50257)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--trace", action="store_true",
                    help="wait accounting per CTA (needs B200_GEMM_TRACE=1 python -m csrc.build)")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    M, D, V = a.rows, a.dim, a.vocab
    ld = (V + 127) // 128 * 128
    x = (torch.randn(M, D, device=dev) * 1.0).to(torch.bfloat16)
    w = (torch.randn(V, D, device=dev) * 0.02).to(torch.bfloat16)
    t = torch.randint(0, V, (M,), device=dev)
    out = torch.empty((M, ld), dtype=torch.bfloat16, device=dev)
    st = stream()
    stats = torch.empty(4, dtype=torch.float32, device=dev)
    lmstate = torch.zeros(8, dtype=torch.float32, device=dev)
    xnorm = torch.empty(M, dtype=torch.float32, device=dev)
    call("b200_lmhead_begin", t.data_ptr(), stats.data_ptr(), lmstate.data_ptr(), -100, M, st)
    call("b200_lmhead_norms", x.data_ptr(), x.stride(0), w.data_ptr(), w.stride(0), M, V, D, xnorm.data_ptr(),
         lmstate.data_ptr(), st)
    max_parts = (V + 127) // 128 + 2
    part = torch.empty((max_parts, M, 2), dtype=torch.float32, device=dev)
    tgt_val = torch.empty(M, dtype=torch.float32, device=dev)
    nparts = ctypes.c_int(0)

    def lm(store):
        call("b200_lmhead_gemm_bf16", x.data_ptr(), w.data_ptr(), out.data_ptr() if store else 0, M, V, D, x.stride(0),
             w.stride(0), ld, xnorm.data_ptr(), lmstate.data_ptr(), part.data_ptr(), max_parts, ctypes.addressof(nparts),
             t.data_ptr(), tgt_val.data_ptr(), 0, st)

    res = {"rows": M, "dim": D, "vocab": V, "staging2": os.environ.get("B200GPT_GEMM_STAGING2", "1")}
    for name, fn in (("plain_logits_bf16", lambda: gemm(x, w, out, M, V, D)),
                     ("lm_softmax_epilogue_store", lambda: lm(True)),
                     ("lm_softmax_epilogue_stats_only", lambda: lm(False))):
        med, best = timed(fn, a.iters)
        res[name] = {"median_us": round(med, 1), "best_us": round(best, 1),
                     "write_TBps": round(M * V * 2 / med / 1e6, 2) if "stats_only" not in name else 0.0,
                     "TFLOPs": round(2.0 * M * V * D / med / 1e6, 1)}
    if a.trace:
        dbg = torch.zeros(4096 * 32, dtype=torch.int64, device=dev)
        for name, fn in (("plain_logits_bf16", lambda: gemm(x, w, out, M, V, D)),
                         ("lm_softmax_epilogue_store", lambda: lm(True)),
                         ("lm_softmax_epilogue_stats_only", lambda: lm(False))):
            dbg.zero_()
            torch.cuda.synchronize()
            call("b200_gemm_set_debug_buffer", dbg.data_ptr())
            fn()
            torch.cuda.synchronize()
            call("b200_gemm_set_debug_buffer", None)
            d = dbg.view(-1, 32).cpu()
            used = d[d[:, 0] > 0]
            lead = used[(used[:, 14] + used[:, 15]) > 0]
            med = lambda t: float(t.float().median())  # noqa: E731
            res[name]["trace_cycles_per_cta"] = {
                "cta_life": med(used[:, 9] - used[:, 0]), "mma_waits_operands": med(lead[:, 14]),
                "mma_waits_accumulator_drained": med(lead[:, 15]), "epilogue_waits_mma": med(used[:, 16]),
                "epilogue_waits_staging": med(used[:, 18]), "producer_waits_free_slot": med(used[:, 17])}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
