"""Build libpgk_b200.so in-tree with nvcc for sm_100a (no torch, no cmake).

    python -m pygpukit_b200.build          # incremental
    python -m pygpukit_b200.build --force

The shared library is the product's only native artefact; the Python package loads it with
ctypes (pygpukit_b200/_lib.py) and fails loudly when it is missing.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB = os.path.join(HERE, "libpgk_b200.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC"]

# quant_nvf4.cu must use the reference's fast-math flags to stay bit-exact (see its header)
SOURCES = {
    "runtime.cu": [],
    "gemv_fp8.cu": [],
    "gemv_nvf4.cu": [],
    "gemv_bf16.cu": [],
    "gemm_w8a16.cu": [],
    "gemm_tc.cu": [],
    "gemm_tc_nvf4.cu": [],
    "gemm_tc_f8.cu": [],
    "elementwise.cu": [],
    "attention.cu": [],
    "moe.cu": [],
    "sampling.cu": [],
    "comm.cu": [],
    "p2p.cu": [],
    "step.cu": [],
    "init.cu": [],
    "quant_nvf4.cu": ["--use_fast_math"],
}


def _deps_mtime() -> float:
    t = 0.0
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for f in os.listdir(root):
            if f.endswith((".cuh", ".h")):
                t = max(t, os.path.getmtime(os.path.join(root, f)))
    return t


def _compile(src: str, extra: list[str], force: bool, hdr_t: float) -> str:
    obj = os.path.join(OBJ, src.replace(".cu", ".o"))
    path = os.path.join(CSRC, src)
    if not force and os.path.exists(obj) and os.path.getmtime(obj) >= max(os.path.getmtime(path), hdr_t):
        return obj
    cmd = [NVCC, *ARCH, *COMMON, *extra, "-c", path, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj


def build(force: bool = False, verbose: bool = True) -> str:
    os.makedirs(OBJ, exist_ok=True)
    hdr_t = _deps_mtime()
    srcs = {s: e for s, e in SOURCES.items() if os.path.exists(os.path.join(CSRC, s))}
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda kv: _compile(kv[0], kv[1], force, hdr_t), srcs.items()))
    newest = max(os.path.getmtime(o) for o in objs)
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < newest:
        cmd = [NVCC, *ARCH, "-shared", "-o", LIB, *objs, "-ldl", "-Xlinker", "--no-undefined"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        if verbose:
            print(f"built {LIB}")
    elif verbose:
        print(f"up to date: {LIB}")
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv)
