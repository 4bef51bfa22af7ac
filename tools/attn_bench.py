"""Attention kernel timings (CUDA events, L2 flushed between iterations) at the shapes of the two model configs:
forward per-tile vs warp-specialised, backward; TFLOP/s counted on the causal half (4 * B * H * T^2/2 * D forward,
2.5x that backward)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402

from distributed_pytorch_cookbook_b200.ops import _lib  # noqa: E402

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


for (B, T, H, D) in [(8, 1023, 16, 128), (64, 255, 8, 32), (4, 4096, 16, 128), (8, 1024, 16, 64)]:
    qkv = (torch.randn(B, T, 3 * H * D, device="cuda") * 0.7).to(torch.bfloat16)
    out = torch.empty((B, T, H * D), dtype=torch.bfloat16, device="cuda")
    lse = torch.empty((B, H, T), dtype=torch.float32, device="cuda")
    dout = torch.randn_like(out)
    dqkv = torch.empty_like(qkv)
    delta = torch.empty((B, H, T), dtype=torch.float32, device="cuda")
    dq_acc = torch.empty((B, T, H * D), dtype=torch.float32, device="cuda")
    scale = D ** -0.5
    flops = 4.0 * B * H * (T * T / 2) * D
    res = {}
    for ws in (0, 1):
        res[ws] = timeit(lambda: _lib.call("b200_attn_fwd_select", qkv.data_ptr(), out.data_ptr(), lse.data_ptr(), B, T, H,
                                           D, scale, None, ws, _lib.stream()))
    bw = {}
    for mode in (0, 1, 2):
        bw[mode] = timeit(lambda: _lib.call("b200_attn_bwd_select", qkv.data_ptr(), out.data_ptr(), dout.data_ptr(),
                                            lse.data_ptr(), dqkv.data_ptr(), delta.data_ptr(), dq_acc.data_ptr(), B, T, H,
                                            D, scale, None, mode, _lib.stream()))
    print(f"B{B} T{T} H{H} D{D}: fwd per-tile {res[0]:7.1f} us ({flops/res[0]/1e6:5.0f} TF)  warp-spec {res[1]:7.1f} us "
          f"({flops/res[1]/1e6:5.0f} TF) | bwd per-tile {bw[0]:7.1f} us ({2.5*flops/bw[0]/1e6:5.0f} TF)  warp-spec+red "
          f"{bw[1]:7.1f} us ({2.5*flops/bw[1]/1e6:5.0f} TF)  warp-spec+dQ kernel {bw[2]:7.1f} us ({2.5*flops/bw[2]/1e6:5.0f} TF)",
          flush=True)
