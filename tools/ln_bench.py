"""Time the LayerNorm backward kernels (CUDA events, 1 GPU): plain, and with the producer hand-off
(dz = dx * relu'(mask) as bf16 + bias-gradient column sums, b200_layernorm_bwd_fused).

    python tools/ln_bench.py [--rows 16320] [--dim 256]        # B200GPT_LN_ROWS=1|2|3|4 picks the rows per warp iteration

Four rotating buffer sets (x, dy, dres, dx, dz, mask: ~75 MB each at the default shape) keep the working set
larger than L2.  For comparison the separate pass the hand-off removes (b200_act_bwd on the fp32 dx) is timed too.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_pytorch_cookbook_b200.ops.native import call, colsum_scratch, stream  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=16320)
    ap.add_argument("--dim", type=int, default=256)
    ap.add_argument("--iters", type=int, default=40)
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    R, D = a.rows, a.dim
    sets = []
    for _ in range(4):
        x = torch.randn(R, D, device=dev)
        sets.append(dict(x=x, dy=torch.randn(R, D, device=dev).bfloat16(), dres=torch.randn(R, D, device=dev),
                         mean=x.mean(1), rstd=1.0 / x.std(1), dx=torch.empty(R, D, device=dev),
                         dz=torch.empty(R, D, device=dev, dtype=torch.bfloat16),
                         mask=torch.randn(R, D, device=dev).bfloat16()))
    gamma = torch.ones(D, device=dev)
    dg, db, dbz = (torch.zeros(D, device=dev) for _ in range(3))
    sc, cnt = colsum_scratch(dev)
    st = stream()

    def ln(s, fuse, mask):
        call("b200_layernorm_bwd_fused", s["dy"].data_ptr(), s["x"].data_ptr(), gamma.data_ptr(), s["mean"].data_ptr(),
             s["rstd"].data_ptr(), s["dres"].data_ptr(), s["dx"].data_ptr(), s["dz"].data_ptr() if fuse else None,
             dg.data_ptr(), db.data_ptr(), R, D, sc.data_ptr(), cnt.data_ptr(), s["mask"].data_ptr() if mask else None,
             1 if mask else 0, dbz.data_ptr() if fuse else None, int(fuse), st)

    def act_bwd(s):
        call("b200_act_bwd", s["dx"].data_ptr(), s["mask"].data_ptr(), s["dz"].data_ptr(), dbz.data_ptr(), R, D, 1, 1,
             sc.data_ptr(), cnt.data_ptr(), st)

    def timed(fn):
        for i in range(4):
            fn(sets[i])
        torch.cuda.synchronize()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(a.iters + 1)]
        ev[0].record()
        for i in range(a.iters):
            fn(sets[i % 4])
            ev[i + 1].record()
        torch.cuda.synchronize()
        ts = sorted(ev[i].elapsed_time(ev[i + 1]) * 1e3 for i in range(a.iters))
        return round(ts[len(ts) // 2], 2)

    res = {"rows": R, "dim": D, "rows_per_iter": os.environ.get("B200GPT_LN_ROWS", "1"),
           "ln_bwd_us": timed(lambda s: ln(s, False, False)),
           "ln_bwd_handoff_cast_us": timed(lambda s: ln(s, True, False)),
           "ln_bwd_handoff_relu_mask_us": timed(lambda s: ln(s, True, True)),
           "separate_act_bwd_us": timed(act_bwd)}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
