"""Build `libovmono3d_lift.so` (the C-ABI library of include/ovmono3d_lift.h) in-tree with nvcc for sm_100a.

    python -m ovmono3d_b200.build [--force]

nvcc cross-compiles without a GPU.  Objects go to `ovmono3d_b200/csrc/build/`, the library next to this file so
that it travels with the source tree (it is git-ignored, not gpurun-ignored).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from typing import List

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(CSRC, "build")
LIB_PATH = os.path.join(HERE, "libovmono3d_lift.so")
SOURCES = ["roi_align.cu", "gemm_tcgen05.cu", "head_decode.cu", "fc_tail.cu", "detect_post.cu", "iou3d.cu", "records.cu", "lift_api.cu"]
HEADERS = [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "decode_core.cuh"), os.path.join(os.path.dirname(HERE), "include", "ovmono3d_lift.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--fmad=true"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the lift has no CPU fallback and cannot be built without the CUDA toolkit")


def _stale(target: str, deps: List[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd: List[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
    if r.stderr.strip():
        sys.stderr.write(r.stderr)


def build_library(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(BUILD, exist_ok=True)
    nvcc = _nvcc()
    objs, jobs = [], []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(BUILD, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + HEADERS):
            extra = ["-Xptxas", "-v"] if verbose else []
            jobs.append([nvcc] + NVCC_FLAGS + extra + ["-c", s, "-o", o])
    with ThreadPoolExecutor(max_workers=max(1, len(jobs))) as ex:
        list(ex.map(_run, jobs))
    if force or jobs or _stale(LIB_PATH, objs):
        _run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB_PATH] + objs)
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
