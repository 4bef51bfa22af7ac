"""Split-K / tile sweep for the small weight-gradient GEMMs of the default GPT (dW[N_out, K_in] += dZ^T X over 16320
tokens, both operands MN-major, fp32 TMA reduce-add): time per launch for every (tile width, CTA pair, splits).

    python tools/wgrad_sweep.py
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_pytorch_cookbook_b200.ops import _lib  # noqa: E402
from distributed_pytorch_cookbook_b200.ops.native import stream  # noqa: E402

dev = torch.device("cuda:0")
TOK = 16320
SHAPES = [(768, 256), (1024, 256), (256, 1024), (256, 256)]     # (rows of dW, cols of dW)


def timed(fn, iters=30):
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
    return round(ts[len(ts) // 2], 2)


for (n_out, k_in) in SHAPES:
    # four rotating operand sets (about 4 x 40 MB) so the operands are not simply L2-resident
    sets = [(torch.randn(TOK, n_out, device=dev).bfloat16(), torch.randn(TOK, k_in, device=dev).bfloat16())
            for _ in range(4)]
    out = torch.zeros(n_out, k_in, device=dev)
    st = stream()
    res = {"dW": [n_out, k_in], "tokens": TOK}
    for bn, pair in ((256, 1), (256, 0), (128, 0)):
        if k_in < bn:
            continue
        for splits in (0, 4, 8, 12, 16, 24, 32, 48):
            it = [0]

            def run():
                dz, x = sets[it[0] % 4]
                it[0] += 1
                _lib.call("b200_gemm_debug", dz.data_ptr(), x.data_ptr(), out.data_ptr(), n_out, k_in, TOK, dz.stride(0),
                          x.stride(0), out.stride(0), 1, 1, -1, -1, -1, -1, -1, -1, bn | (pair << 17),
                          splits if splits else 2, st)
            if splits == 0:
                continue
            try:
                res[f"bn{bn}_{'pair' if pair else 'single'}_s{splits}"] = timed(run)
            except Exception as e:  # noqa: BLE001
                res[f"bn{bn}_{'pair' if pair else 'single'}_s{splits}"] = str(e)[:60]
    print(json.dumps(res))
