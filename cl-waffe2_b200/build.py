"""Build libwaffe2cuda.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python cl-waffe2_b200/build.py [--force]

Objects go to cl-waffe2_b200/_obj/, the library to cl-waffe2_b200/lib/libwaffe2cuda.so (git-ignored,
but it travels to the GPU box with the gpurun snapshot)."""
from __future__ import annotations

import concurrent.futures as cf
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "_obj")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libwaffe2cuda.so")
SOURCES = ["runtime", "elementwise", "reduce", "softmax", "misc", "fused", "comm", "p2p", "conv", "gemm", "gemm_tc", "attention"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def _newer(src_list, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in src_list)


def _compile(name: str, force: bool) -> str:
    src = os.path.join(CSRC, name + ".cu")
    obj = os.path.join(OBJ, name + ".o")
    deps = [src] + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh")] + \
           [os.path.join(HERE, "..", "include", "waffe2cuda.h")]
    if not force and not _newer(deps, obj):
        return obj
    r = subprocess.run([NVCC] + FLAGS + ["-c", src, "-o", obj], capture_output=True, text=True)
    with open(os.path.join(OBJ, name + ".log"), "w") as f:
        f.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {name}.cu:\n{r.stdout}\n{r.stderr}")
    return obj


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    with cf.ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        objs = list(ex.map(lambda n: _compile(n, force), SOURCES))
    if force or _newer(objs, LIB):
        r = subprocess.run([NVCC, "-shared", "-o", LIB] + objs + ["-ldl"], capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    if verbose:
        print("built", LIB)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
