"""CTA-pair (tcgen05 cta_group::2) GEMM bring-up: numerics against torch fp32 and timing against the
single-CTA kernel on the framework's real shapes.  One B200, run under its own `timeout`.

    python tools/gemm_2sm_check.py [--iters 30]
"""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402

from distributed_pytorch_cookbook_b200.ops import _lib, native  # noqa: E402

ONE, TWO = 0x800, 0x400     # cfg bits: force single CTA / force CTA pair


def ceil(a, b):
    return (a + b - 1) // b * b


def make(rows, cols, dev):
    """bf16 [rows, cols] view with a 16-byte aligned row stride."""
    ld = ceil(cols, 8)
    t = (torch.randn(rows, ld, device=dev) * 0.5).to(torch.bfloat16)
    return t[:, :cols]


def run_case(name, M, N, K, a_mn, b_mn, epi, iters, dev, order=0):
    A = make(K, M, dev) if a_mn else make(M, K, dev)
    B = make(K, N, dev) if b_mn else make(N, K, dev)
    Af = (A.t() if a_mn else A).float()
    Bf = (B.t() if b_mn else B).float()
    ref = Af @ Bf.t()
    kw = {}
    out_dtype = torch.bfloat16
    base = None
    if epi == "acc":            # fp32 accumulate (wgrad / split-K dgrad)
        out_dtype = torch.float32
        base = torch.randn(M, N, device=dev)
        ref = ref + base
        kw["accumulate"] = True
    elif epi == "bias_relu":
        bias = torch.randn(N, device=dev)
        ref = torch.relu(ref + bias)
        kw.update(bias=bias, act=1)
    elif epi == "res":          # bias + fp32 residual -> fp32 out
        out_dtype = torch.float32
        bias = torch.randn(N, device=dev)
        res = torch.randn(M, N, device=dev)
        ref = ref + bias + res
        kw.update(bias=bias, residual=res)
    elif epi == "mask":
        msk = make(M, N, dev)
        ref = torch.where(msk.float() > 0, ref, torch.zeros_like(ref))
        kw["mask_src"] = msk
    ldo = ceil(N, 8 if out_dtype == torch.bfloat16 else 4)
    res_line = {}
    for tag, bit in (("1sm", ONE | order), ("2sm", TWO | order)):
        store = torch.zeros(M, ldo, device=dev, dtype=out_dtype)
        out = store[:, :N]
        if base is not None:
            out.copy_(base)
        native.gemm(A, B, out, M, N, K, a_mn_major=a_mn, b_mn_major=b_mn, cfg=bit, **kw)
        torch.cuda.synchronize()
        import ctypes
        ew = ctypes.c_int(0)
        _lib.call("b200_device_error", ctypes.addressof(ew), 1)
        err = ew.value
        rel = float((out.float() - ref).norm() / ref.norm())
        # timing (accumulate keeps accumulating: harmless for time)
        for _ in range(3):
            native.gemm(A, B, out, M, N, K, a_mn_major=a_mn, b_mn_major=b_mn, cfg=bit, **kw)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(iters):
            native.gemm(A, B, out, M, N, K, a_mn_major=a_mn, b_mn_major=b_mn, cfg=bit, **kw)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / iters
        res_line[tag] = (rel, us, err)
    tf = 2.0 * M * N * K / 1e12
    r1, r2 = res_line["1sm"], res_line["2sm"]
    ok = r2[0] < 1.5e-2 and r2[2] == 0
    print(f"{name:28s} M={M:6d} N={N:6d} K={K:6d} a_mn={int(a_mn)} b_mn={int(b_mn)} epi={epi:9s} "
          f"rel 1sm={r1[0]:.2e} 2sm={r2[0]:.2e} err={r2[2]}  "
          f"us 1sm={r1[1]:7.1f} ({tf / r1[1] * 1e6:6.0f} TF) 2sm={r2[1]:7.1f} ({tf / r2[1] * 1e6:6.0f} TF)  "
          f"{'OK' if ok else 'FAIL'}", flush=True)
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--quick", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    T = 16320
    cases = [
        ("small K-major", 512, 256, 128, False, False, "none"),
        ("small tails", 300, 328, 200, False, False, "none"),
        ("small b_mn", 512, 512, 256, False, True, "none"),
        ("small a_mn b_mn acc", 512, 256, 512, True, True, "acc"),
    ]
    if not a.quick:
        cases += [
            ("lm fwd", 4096, 50257, 256, False, False, "none"),
            ("lm fwd n-fastest", 4096, 50257, 256, False, False, "none", 0x1000),
            ("lm wgrad", 50257, 256, 4096, True, True, "acc"),
            ("lm dgrad", 4096, 256, 50257, False, True, "acc"),
            ("fwd qkv", T, 768, 256, False, False, "none"),
            ("fwd up", T, 1024, 256, False, False, "bias_relu"),
            ("fwd out-proj", T, 256, 256, False, False, "res"),
            ("fwd down", T, 256, 1024, False, False, "res"),
            ("dgrad qkv", T, 256, 768, False, True, "none"),
            ("dgrad down(mask)", T, 1024, 256, False, True, "mask"),
            ("wgrad up", 1024, 256, T, True, True, "acc"),
            ("wgrad qkv", 768, 256, T, True, True, "acc"),
            ("1.3b fwd up", 8184, 8192, 2048, False, False, "bias_relu"),
            ("1.3b fwd up n-fastest", 8184, 8192, 2048, False, False, "bias_relu", 0x1000),
            ("1.3b dgrad", 8184, 2048, 8192, False, True, "none"),
            ("1.3b wgrad", 8192, 2048, 8184, True, True, "acc"),
        ]
    # context: pure write and read+write bandwidth of this GPU (1 GiB buffers)
    big = torch.empty(1 << 28, dtype=torch.float32, device=dev)
    big2 = torch.empty_like(big)
    for name, fn, nbytes in (("fill (write only)", lambda: big.fill_(1.0), big.numel() * 4),
                             ("copy (read + write)", lambda: big2.copy_(big), big.numel() * 8)):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn()
        e1.record()
        torch.cuda.synchronize()
        print(f"{name}: {nbytes * 10 / (e0.elapsed_time(e1) * 1e-3) / 1e12:.2f} TB/s", flush=True)
    del big, big2
    bad = 0
    for c in cases:
        try:
            bad += 0 if run_case(*c[:7], a.iters, dev, *(c[7:])) else 1
        except Exception as e:  # noqa: BLE001
            bad += 1
            print(f"{c[0]:28s} EXCEPTION {e}", flush=True)
    print("failures:", bad)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
