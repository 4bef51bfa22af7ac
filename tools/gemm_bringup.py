"""GEMM bring-up on a real B200: checks the tcgen05 kernel against torch.matmul (fp32 reference
of the bf16 inputs) for every operand-major combination, and - if a combination fails with the
default UMMA descriptor encoding - sweeps alternative LBO/SBO/K-step encodings through
b200_gemm_debug so that one GPU session is enough to pin the right one.  Writes JSON."""
import itertools
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch

from distributed_pytorch_cookbook_b200.ops import _lib


def dev_err():
    import ctypes
    v = ctypes.c_int(0)
    _lib.call("b200_device_error", ctypes.addressof(v), 1)
    return v.value


def run_case(M, N, K, a_mn, b_mn, ov=None, bn=0, splits=0):
    torch.manual_seed(0)
    Kp = (K + 7) // 8 * 8  # 16-byte row strides for TMA
    A = torch.zeros(M, Kp, device="cuda", dtype=torch.bfloat16)
    B = torch.zeros(N, Kp, device="cuda", dtype=torch.bfloat16)
    A[:, :K] = torch.randn(M, K, device="cuda").to(torch.bfloat16)
    B[:, :K] = torch.randn(N, K, device="cuda").to(torch.bfloat16)
    ref = A.float() @ B.float().t()
    Am = A.t().contiguous() if a_mn else A
    Bm = B.t().contiguous() if b_mn else B
    C = torch.zeros(M, N, device="cuda", dtype=torch.float32)
    o = ov or (-1,) * 6
    _lib.call("b200_gemm_debug", Am.data_ptr(), Bm.data_ptr(), C.data_ptr(), M, N, K,
              Am.stride(0), Bm.stride(0), C.stride(0), int(a_mn), int(b_mn), *o, bn, splits,
              _lib.stream())
    torch.cuda.synchronize()
    e = dev_err()
    err = (C - ref).abs().max().item()
    scale = ref.abs().max().item()
    return {"M": M, "N": N, "K": K, "a_mn": a_mn, "b_mn": b_mn, "ov": o, "bn": bn, "splits": splits,
            "max_err": err, "ref_max": scale, "dev_err": e, "ok": bool(err <= 2e-2 * max(scale, 1.0) and e == 0)}


def main():
    out = {"cases": [], "sweeps": {}}
    shapes = [(128, 128, 64), (256, 256, 256), (16320, 768, 256), (1000, 333 * 8, 1024), (300, 50257, 256)]
    for a_mn, b_mn in [(0, 0), (0, 1), (1, 0), (1, 1)]:
        for (M, N, K) in shapes:
            if (a_mn and M % 8) or (b_mn and N % 8):
                continue
            for bn in (128, 256):
                r = run_case(M, N, K, a_mn, b_mn, bn=bn)
                out["cases"].append(r)
                print(r, flush=True)
    # split-K
    for r in [run_case(256, 256, 16320, 1, 1, splits=9), run_case(1024, 256, 50257, 0, 1, splits=16)]:
        out["cases"].append(r)
        print(r, flush=True)
    # sweep alternatives for failing major combos
    for a_mn, b_mn in [(0, 1), (1, 0), (1, 1), (0, 0)]:
        base = [c for c in out["cases"] if c["a_mn"] == a_mn and c["b_mn"] == b_mn and c["splits"] == 0]
        if all(c["ok"] for c in base):
            continue
        res = []
        mn_opts = [(8192, 1024, 2048), (1024, 8192, 2048), (8192, 1024, 4096), (128, 1024, 2048),
                   (1024, 128, 2048), (8192, 128, 256), (16, 1024, 2048), (8192, 1024, 1024)]
        k_opts = [(16, 1024, 32), (0, 1024, 32), (1024, 1024, 32), (16, 1024, 16)]
        for oa, ob in itertools.product(mn_opts if a_mn else k_opts, mn_opts if b_mn else k_opts):
            r = run_case(256, 256, 256, a_mn, b_mn, ov=(*oa, *ob), bn=128)
            res.append(r)
            if r["ok"]:
                print("SWEEP HIT", r, flush=True)
        out["sweeps"][f"{a_mn}{b_mn}"] = res
    path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/gemm_bringup.json"
    json.dump(out, open(path, "w"), indent=1)
    n_ok = sum(c["ok"] for c in out["cases"])
    print(f"GEMM bring-up: {n_ok}/{len(out['cases'])} cases ok")


if __name__ == "__main__":
    main()
