"""ctypes binding to the in-tree sm_100a library (``_C/libb200gpt.so``).

The library is plain C ABI (``include/b200gpt.h``): raw device pointers, sizes and a
``cudaStream_t``; nothing in it depends on ATen, so it builds in seconds with nvcc
(``python -m csrc.build``) and the built ``.so`` ships to the GPU box in-tree.  Every
entry point returns 0 on success or a negative error code; ``call`` turns a non-zero
return into a Python exception carrying ``b200_last_error()``.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_float, c_int, c_int64, c_void_p
from pathlib import Path
from typing import Optional

import torch

_SO = Path(__file__).resolve().parent.parent / "_C" / "libb200gpt.so"
_lib: Optional[ctypes.CDLL] = None
_err: Optional[str] = None
_tried = False

P, I, L, Fl = c_void_p, c_int, c_int64, c_float

# name -> argtypes (return type is always int)
_SIGS = {
    "b200_embedding_fwd": [P, P, P, P, P, I, I, I, I, P],
    "b200_embedding_bwd": [P, P, P, P, P, I, I, I, I, P],
    "b200_layernorm_fwd": [P, P, P, P, P, P, I, I, Fl, P],
    "b200_layernorm_bwd": [P, P, P, P, P, P, P, P, P, P, I, I, P, P, P],
    "b200_layernorm_bwd_fused": [P, P, P, P, P, P, P, P, P, P, I, I, P, P, P, I, P, I, P],
    "b200_adamw": [P, P, P, P, P, L, P, Fl, Fl, Fl, Fl, P],
    "b200_adamw_tick": [P, P],
    "b200_cast_bf16": [P, P, L, P],
    "b200_relu_bwd": [P, P, P, P, I, I, I, P, P, P],
    "b200_act_bwd": [P, P, P, P, I, I, I, I, P, P, P],
    "b200_gemm_dact_colsum_bf16": [P, P, P, I, I, I, I, I, I, I, P, P, I, I, P],
    "b200_colsum_bf16": [P, P, I, I, I, P, P, P],
    "b200_relu_bwd_static": [P, P, P, P, I, I, I, P],
    "b200_count_valid": [P, P, I, L, P],
    "b200_device_error": [P, I],
    "b200_ce_fwd_bwd": [P, P, P, I, I, I, I, I, Fl, P],
    "b200_lmhead_begin": [P, P, P, I, L, P],
    "b200_lmhead_norms": [P, I, P, I, I, I, I, P, P, P],
    "b200_lmhead_gemm_bf16": [P, P, P, I, I, I, I, I, I, P, P, P, I, P, P, P, I, P],
    "b200_lmhead_finish": [P, I, I, P, P, P, I, P, P, P, I, I, P, I, I, P, I, P, P, Fl, I, P],
    "b200_gemm_rowscale_bf16": [P, P, P, I, I, I, I, I, I, I, P, I, P],
    "b200_gemm_bf16": [P, P, P, I, I, I, I, I, I, I, I, P, P, I, I, I, I, P, P, I, P],
    "b200_gemm_debug": [P, P, P] + [I] * 16 + [P],
    "b200_gemm_wgrad_group": [I, P, P, P, P, P, I, P, P, P, P],
    "b200_gemm_mask_colsum_bf16": [P, P, P, I, I, I, I, I, I, I, P, P, I, P],
    "b200_attn_fwd": [P, P, P, I, I, I, I, Fl, P, P],
    "b200_attn_fwd_select": [P, P, P, I, I, I, I, Fl, P, I, P],
    "b200_attn_bwd": [P, P, P, P, P, P, P, I, I, I, I, Fl, P, P],
    "b200_attn_bwd_select": [P, P, P, P, P, P, P, I, I, I, I, Fl, P, I, P],
    "b200_allreduce": [P, P, P, I, I, L, L, Fl, I, I, P],
    "b200_allreduce_bf16": [P, P, P, P, I, I, L, L, Fl, I, I, P],
    "b200_rs_adamw": [P, P, P, I, I, L, L, Fl, P, P, P, P, P, Fl, Fl, Fl, Fl, I, I, I, P],
    "b200_reduce_scatter": [P, P, P, I, I, L, L, Fl, P, I, I, I, P],
    "b200_allgather": [P, P, I, I, L, L, P, I, I, I, P],
    "b200_ag_gemm_bf16": [P, P, P, I, I, I, I, I, I, P, P, I, I, I, P, P, P, L, L, L, P, I, P],
    "b200_spin": [L, P],
    "b200_set_comm_timeout_ms": [I],
    "b200_barrier": [P, I, I, I, P],
    "b200_smoke": [P],
    "b200_gemm_set_debug_buffer": [P],
    "b200_attn_set_debug_buffer": [P],
}


def so_path() -> Path:
    return _SO


def _load():
    global _lib, _err, _tried
    if _tried:
        return
    _tried = True
    if os.environ.get("B200GPT_DISABLE_NATIVE") == "1":
        _err = "disabled by B200GPT_DISABLE_NATIVE=1"
        return
    if not _SO.exists():
        _err = f"{_SO} not built"
        return
    try:
        lib = ctypes.CDLL(str(_SO))
    except OSError as e:  # pragma: no cover - depends on the box
        _err = f"dlopen failed: {e}"
        return
    lib.b200_last_error.restype = ctypes.c_char_p
    lib.b200_last_error.argtypes = []
    for name, argtypes in _SIGS.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.restype = c_int
        fn.argtypes = argtypes
    _lib = lib


def loaded() -> bool:
    """True when the shared object could be dlopen'ed (does not need a GPU)."""
    _load()
    return _lib is not None


def available() -> bool:
    """Library loaded *and* a CUDA device is present."""
    return loaded() and torch.cuda.is_available()


def load_error() -> Optional[str]:
    _load()
    return _err


def has(name: str) -> bool:
    _load()
    return _lib is not None and hasattr(_lib, name)


def ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


_launches = 0


def launch_count() -> int:
    """Number of native entry-point calls made so far (bench.py's ``gpu_launches``)."""
    return _launches


def call(name: str, *args):
    global _launches
    _load()
    if _lib is None:
        raise RuntimeError(f"native library unavailable: {_err}")
    rc = getattr(_lib, name)(*args)
    _launches += 1
    if rc != 0:
        msg = _lib.b200_last_error()
        raise RuntimeError(f"{name} failed rc={rc}: {msg.decode() if msg else ''}")
    return rc
