"""Tile-configuration sweep for the layer GEMMs of the default GPT (16320 tokens, dim 256): CTA-pair 256 x 256 tiles
(the default), single-CTA 128 x 256 tiles, and 128 x 128 tiles with two CTAs per SM.

    python tools/gemm_cfg_sweep.py
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_pytorch_cookbook_b200.ops.native import gemm  # noqa: E402

dev = torch.device("cuda:0")
BF = torch.bfloat16
TOK = 16320
CFGS = {"pair_256x256": 0x400 | 256, "single_128x256": 0x800 | 256, "single_128x128_2cta": 0x800 | 128}


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


def case(name, N, K, **kw):
    res = {"gemm": name, "M": TOK, "N": N, "K": K}
    b_mn = kw.pop("b_mn", False)
    sets = []
    for _ in range(3):
        a = torch.randn(TOK, K, device=dev).to(BF)
        b = (torch.randn(K, N, device=dev) if b_mn else torch.randn(N, K, device=dev)).to(BF)
        sets.append((a, b))
    f32 = kw.pop("f32", False)
    out = torch.empty(TOK, N, device=dev, dtype=torch.float32 if f32 else BF)
    extra = {}
    if kw.get("bias"):
        extra["bias"] = torch.randn(N, device=dev)
    if kw.get("res"):
        extra["residual"] = torch.randn(TOK, N, device=dev)
    if kw.get("aux"):
        extra["aux_out"] = torch.empty(TOK, N, device=dev, dtype=BF)
    if kw.get("act"):
        extra["act"] = 1
    for cname, cfg in CFGS.items():
        it = [0]

        def run():
            a, b = sets[it[0] % 3]
            it[0] += 1
            gemm(a, b, out, TOK, N, K, b_mn_major=b_mn, cfg=cfg, **extra)
        try:
            res[cname] = timed(run)
        except Exception as e:  # noqa: BLE001
            res[cname] = str(e)[:80]
    print(json.dumps(res), flush=True)


case("qkv fwd (plain)", 768, 256)
case("out-proj fwd (+bias +residual, fp32)", 256, 256, bias=True, res=True, f32=True)
case("mlp up fwd (+bias relu)", 1024, 256, bias=True, act=True)
case("mlp down fwd (+bias relu +residual +aux, fp32)", 256, 1024, bias=True, act=True, res=True, aux=True, f32=True)
case("qkv dgrad (W mn-major)", 256, 768, b_mn=True)
case("out-proj / up dgrad", 256, 256, b_mn=True)
case("up-proj dgrad (K = 1024)", 256, 1024, b_mn=True)
