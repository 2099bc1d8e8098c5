"""Build the native library in-tree with nvcc for sm_100a (no GPU needed: nvcc cross-compiles).

    python -m underwater_image_enhancement_b200.build [--force]

Outputs (git-ignored, shipped to the GPU box by gpurun):
    underwater_image_enhancement_b200/libuwie_b200.so   the product: CUDA kernels + C ABI (include/uwie.h)
    tests/_hostcheck.so                                 TEST-ONLY host build of the shared integer arithmetic
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
OBJ = os.path.join(CSRC, "build")
LIB = os.path.join(PKG, "libuwie_b200.so")
HOSTCHECK = os.path.join(ROOT, "tests", "_hostcheck.so")

CU_SOURCES = ("k_stats.cu", "k_setup.cu", "k_canny.cu", "k_quadtree.cu", "k_guided.cu", "k_select.cu",
              "k_apply.cu", "k_score.cu", "k_vgg.cu", "k_features.cu", "k_qa.cu", "k_prims.cu", "k_dct.cu", "k_dct_fft.cu", "uwie_engine.cu")
CPP_SOURCES = ("uwie_tables.cpp",)
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off"]


def _newer(target: str, deps) -> bool:
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps)


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hs.append(os.path.join(ROOT, "include", "uwie.h"))
    return hs


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: " + " ".join(cmd) + "\n" + r.stdout + r.stderr)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    headers = _headers()
    jobs = []
    objs = []
    for src in CU_SOURCES + CPP_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.rsplit(".", 1)[0] + ".o")
        objs.append(o)
        if force or not _newer(o, [s] + headers):
            jobs.append([NVCC] + NVCC_FLAGS + ["-c", s, "-o", o])
    if jobs:
        if verbose:
            print(f"[uwie build] compiling {len(jobs)} translation units for sm_100a", file=sys.stderr)
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(_run, jobs))
    if force or jobs or not _newer(LIB, objs):
        _run([NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"])
    hc_srcs = [os.path.join(CSRC, "hostcheck.cpp"), os.path.join(CSRC, "uwie_tables.cpp")]
    if force or not _newer(HOSTCHECK, hc_srcs + headers):
        _run(["g++", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", HOSTCHECK] + hc_srcs)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
