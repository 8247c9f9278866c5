"""Builds libavenir_tree.so (sm_100a kernels + C-ABI + C++ tool layer) and the avenir-tree CLI, in-tree.

nvcc cross-compiles without a GPU.  The built files are git-ignored but travel to the GPU box.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libavenir_tree.so")
CLI = os.path.join(HERE, "avenir-tree")

CUDA_SOURCES = ["avt_engine.cu"]
HOST_SOURCES = ["host/dtb.cpp", "host/cpg.cpp", "host/mop.cpp", "host/crc.cpp", "host/dap.cpp", "host/cce.cpp", "host/cca.cpp", "host/tool.cpp"]
HEADERS = ["avt_kernels.cuh", "avt_scan.cuh", "avt_scan_fast.cuh", "avt_csv.cuh", "avt_predict.cuh", "avt_bagging.cuh", "avt_forest.cuh", "avt_joint.cuh", "avt_sorted.cuh", "avt_device.cuh", "fdlibm_log.cuh", "host/model.hpp", "host/javaemu.hpp", "host/json.hpp",
           "host/tool.hpp", "../../include/avenir_tree.h"]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-fmad=false",                      # Java never fuses multiply-add
              "-Xcompiler", "-fPIC,-ffp-contract=off,-Wall,-Wno-unknown-pragmas", "-diag-suppress", "39"]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libavenir_tree.so cannot be built (there is no CPU build of this library)")


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def build(force: bool = False, verbose: bool = False) -> str:
    srcs = [os.path.join(CSRC, s) for s in CUDA_SOURCES + HOST_SOURCES]
    deps = srcs + [os.path.join(CSRC, h) for h in HEADERS] + [os.path.abspath(__file__)]
    if force or _stale(LIB, deps):
        cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-shared", "-o", LIB] + srcs + ["-lnccl"]
        if verbose:
            print(" ".join(cmd))
        subprocess.run(cmd, check=True)
    cli_src = os.path.join(CSRC, "cli", "avenir_tree_main.cpp")
    if force or _stale(CLI, [cli_src, LIB]):
        cmd = ["g++", "-O2", "-std=c++17", "-o", CLI, cli_src, "-L" + HERE, "-lavenir_tree", "-Wl,-rpath,$ORIGIN"]
        subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
