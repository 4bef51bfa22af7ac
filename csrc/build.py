"""Build the in-tree sm_100a library: ``python -m csrc.build [--force] [-v]``.

nvcc cross-compiles for sm_100a without a GPU (seconds per file).  Objects go to
``build/``; the shared object to ``distributed_pytorch_cookbook_b200/_C/libb200gpt.so``
(git-ignored, but shipped to the GPU box by gpurun).
"""
from __future__ import annotations

import argparse
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CSRC = ROOT / "csrc"
OUT_DIR = ROOT / "distributed_pytorch_cookbook_b200" / "_C"
SO = OUT_DIR / "libb200gpt.so"
OBJ_DIR = ROOT / "build" / "obj"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v"]
if os.environ.get("B200_GEMM_TRACE") == "1":      # bring-up build: GEMM wait accounting (tools/gemm_phases.py)
    FLAGS.append("-DB200_GEMM_TRACE")


def sources():
    return sorted(CSRC.glob("*.cu"))


def _digest(src: Path) -> str:
    h = hashlib.sha256()
    h.update(" ".join(FLAGS).encode())
    for f in [src, *sorted(CSRC.glob("*.cuh")), *sorted((ROOT / "include").glob("*.h"))]:
        h.update(f.read_bytes())
    return h.hexdigest()


def _compile(src: Path, force: bool, verbose: bool):
    obj = OBJ_DIR / (src.stem + ".o")
    stamp = OBJ_DIR / (src.stem + ".sha")
    dig = _digest(src)
    if not force and obj.exists() and stamp.exists() and stamp.read_text() == dig:
        return obj, "", False
    cmd = [NVCC, *FLAGS, "-c", str(src), "-o", str(obj)]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    stamp.write_text(dig)
    (OBJ_DIR / (src.stem + ".ptxas.log")).write_text(r.stderr)
    return obj, r.stderr, True


def build(force: bool = False, verbose: bool = False) -> Path:
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    srcs = sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(lambda s: _compile(s, force, verbose), srcs))
    rebuilt = any(r[2] for r in results)
    if verbose:
        for (_, log, did), s in zip(results, srcs):
            if did:
                print(f"== {s.name}\n{log}")
    if rebuilt or not SO.exists():
        cmd = [NVCC, "-shared", "-o", str(SO), *[str(r[0]) for r in results],
               "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return SO


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("-v", "--verbose", action="store_true")
    a = ap.parse_args()
    so = build(a.force, a.verbose)
    print(so)
