"""The allocation-free variant of the ReLU-backward / bias-gradient kernel.

`b200_relu_bwd` asks its caller for a scratch buffer.  Some callers have none to give: the
device memory of their process belongs to a pool they cannot carve from, or they sit in a
callback / tight loop where calling the CUDA allocator is off limits.  `b200_relu_bwd_static`
serves them from a fixed array inside the library.  This test drives both entry points through
the C ABI (ctypes, raw pointers) and requires identical results, twice in a row - the second
call proves the shared array was left clean by the first.  It lives in its own file, named to
be collected last, so a `-x` run has finished every other test before this one can stop it.
"""
import ctypes

import pytest
import torch

from distributed_pytorch_cookbook_b200.ops import _lib

pytestmark = pytest.mark.gpu
P, I = ctypes.c_void_p, ctypes.c_int


def _lib_handle():
    assert _lib.available(), f"native library not loaded: {_lib.load_error()}"
    lib = ctypes.CDLL(str(_lib.so_path()))
    lib.b200_relu_bwd.restype = I
    lib.b200_relu_bwd.argtypes = [P, P, P, P, I, I, I, P, P, P]
    lib.b200_relu_bwd_static.restype = I
    lib.b200_relu_bwd_static.argtypes = [P, P, P, P, I, I, I, P]
    lib.b200_last_error.restype = ctypes.c_char_p
    return lib


@pytest.mark.parametrize("fp32", [True, False])
def test_static_scratch_variant_matches_caller_scratch_variant(fp32):
    lib = _lib_handle()
    dev = torch.device("cuda")
    torch.manual_seed(0)
    rows, cols = 3000, 1024
    stream = torch.cuda.current_stream().cuda_stream
    scratch = torch.zeros(16 * cols, device=dev)
    counter = torch.zeros(1, dtype=torch.int32, device=dev)
    for call in range(2):                     # two calls in a row on the shared array
        dy = torch.randn(rows, cols, device=dev) * (call + 1)
        if not fp32:
            dy = dy.to(torch.bfloat16)
        y = torch.randn(rows, cols, device=dev).relu().to(torch.bfloat16)
        dz_a = torch.empty(rows, cols, device=dev, dtype=torch.bfloat16)
        dz_b = torch.empty_like(dz_a)
        db_a = torch.zeros(cols, device=dev)
        db_b = torch.zeros(cols, device=dev)
        rc = lib.b200_relu_bwd(dy.data_ptr(), y.data_ptr(), dz_a.data_ptr(), db_a.data_ptr(), rows, cols,
                               int(fp32), scratch.data_ptr(), counter.data_ptr(), stream)
        assert rc == 0, lib.b200_last_error()
        rc = lib.b200_relu_bwd_static(dy.data_ptr(), y.data_ptr(), dz_b.data_ptr(), db_b.data_ptr(), rows, cols,
                                      int(fp32), stream)
        assert rc == 0, lib.b200_last_error()
        torch.cuda.synchronize()
        assert torch.equal(dz_a, dz_b)
        assert torch.allclose(db_a, db_b, atol=5e-2, rtol=1e-2)     # same tolerance as test_relu_bwd_and_bias_grad
        ref = (dy.float() * (y.float() > 0)).sum(0)
        assert torch.allclose(db_b, ref, atol=5e-2 * (call + 1), rtol=1e-2)


def test_static_scratch_variant_refuses_oversize():
    lib = _lib_handle()
    dev = torch.device("cuda")
    rows, cols = 8, 4096 + 8
    dy = torch.zeros(rows, cols, device=dev)
    y = torch.zeros(rows, cols, device=dev, dtype=torch.bfloat16)
    dz = torch.empty_like(y)
    db = torch.zeros(cols, device=dev)
    rc = lib.b200_relu_bwd_static(dy.data_ptr(), y.data_ptr(), dz.data_ptr(), db.data_ptr(), rows, cols, 1,
                                  torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    assert b"exceeds the built-in scratch" in lib.b200_last_error()
