#!/usr/bin/env python
"""Build libb200step.so (the C-ABI library) in-tree with nvcc for sm_100a.

    python kubeflow-v2-distributed-pytorch_b200/build.py [--force] [--verbose]

Objects are compiled in parallel (one nvcc per .cu) into build/, then linked into lib/libb200step.so.
Only sm_100a code is generated: -gencode arch=compute_100a,code=sm_100a (the plain -arch=sm_100a form also
emits a compute_100 pass that rejects tcgen05 PTX).
"""
from __future__ import annotations

import argparse
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(HERE, "build")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libb200step.so")
SOURCES = ["abi.cu", "preprocess.cu", "gemm_tcgen05.cu", "gemm_ce.cu", "loss.cu", "optim.cu", "comm.cu", "conv.cu"]


def _nccl_include() -> str:
    import importlib.util

    spec = importlib.util.find_spec("nvidia.nccl")
    if spec is None or not spec.submodule_search_locations:
        raise RuntimeError("nvidia.nccl (NCCL headers) not found in site-packages")
    return os.path.join(list(spec.submodule_search_locations)[0], "include")


def _nvcc() -> str:
    cand = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    return cand if os.path.exists(cand) else "nvcc"


def _flags(verbose: bool, experiments: bool = False) -> list[str]:
    f = [
        "-std=c++17", "-O3", "-lineinfo",
        "-gencode", "arch=compute_100a,code=sm_100a",
        "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
        "-I", _nccl_include(),
    ]
    if verbose:
        f += ["-Xptxas", "-v"]
    if experiments or os.environ.get("B200STEP_BUILD_EXPERIMENTS") == "1":
        f += ["-DB200STEP_EXPERIMENTS"]   # timing knobs inside the GEMM (results are wrong when they are switched on)
    return f


def _stamp(srcs: list[str], flags: list[str]) -> str:
    h = hashlib.sha256()
    for p in srcs:
        with open(p, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(flags).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False, experiments: bool = False) -> str:
    os.makedirs(BUILD, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    sources = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    flags = _flags(verbose, experiments)
    deps = [os.path.join(CSRC, s) for s in sources] + [
        os.path.join(CSRC, "common.cuh"),
        os.path.join(CSRC, "gemm_common.cuh"),
        os.path.join(HERE, "..", "include", "b200step.h"),
    ]
    stamp = _stamp(deps, flags)
    stamp_file = os.path.join(BUILD, "stamp")
    if not force and os.path.exists(LIB) and os.path.exists(stamp_file):
        with open(stamp_file) as fh:
            if fh.read().strip() == stamp:
                return LIB
    nvcc = _nvcc()

    def compile_one(src: str) -> str:
        obj = os.path.join(BUILD, src.replace(".cu", ".o"))
        cmd = [nvcc, *flags, "-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose or r.returncode != 0:
            sys.stderr.write(f"$ {' '.join(cmd)}\n{r.stdout}{r.stderr}\n")
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}")
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(sources))) as ex:
        objs = list(ex.map(compile_one, sources))
    cmd = [nvcc, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(f"$ {' '.join(cmd)}\n{r.stdout}{r.stderr}\n")
        raise RuntimeError("link failed")
    with open(stamp_file, "w") as fh:
        fh.write(stamp)
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--experiments", action="store_true", help="compile the GEMM timing knobs in (never ship this)")
    a = ap.parse_args()
    print(build(force=a.force, verbose=a.verbose, experiments=a.experiments))
