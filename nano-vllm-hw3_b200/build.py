"""Build libqlinear_b200.so (the C-ABI library of include/qlinear_b200.h) for sm_100a, in-tree.

    python nano-vllm-hw3_b200/build.py [--force]

nvcc cross-compiles without a GPU.  One arch only: -gencode arch=compute_100a,code=sm_100a
(tcgen05 kind::i8 does not exist on any other target).  The result lands next to the Python
binding (nanovllm/_lib/) so it ships with a repository snapshot; it is git-ignored.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT_DIR = os.path.join(HERE, "nanovllm", "_lib")
OUT = os.path.join(OUT_DIR, "libqlinear_b200.so")
OBJ_DIR = os.path.join(HERE, "build")
SOURCES = ["api.cu", "quant.cu", "producers.cu", "allreduce.cu", "tp_collective.cu", "gemv.cu", "gemm_tcgen05.cu", "gemm_2cta.cu", "gemm_skinny.cu", "decode_chain.cu", "static_scratch.cu"]
HEADERS = ["common.cuh", "kernels.h", os.path.join("..", "..", "include", "qlinear_b200.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]


def _sources():
    return [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def _newest_input() -> float:
    files = [os.path.join(CSRC, s) for s in _sources()] + [os.path.join(CSRC, h) for h in HEADERS]
    files.append(os.path.abspath(__file__))
    return max(os.path.getmtime(f) for f in files)


def build_library(force: bool = False, verbose: bool = False) -> str:
    if not force and os.path.exists(OUT) and os.path.getmtime(OUT) >= _newest_input():
        return OUT
    os.makedirs(OBJ_DIR, exist_ok=True)
    os.makedirs(OUT_DIR, exist_ok=True)

    def compile_one(src: str) -> str:
        obj = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
        cmd = [NVCC, *FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, _sources()))
    cmd = [NVCC, "-shared", "-o", OUT, *objs, "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return OUT


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
