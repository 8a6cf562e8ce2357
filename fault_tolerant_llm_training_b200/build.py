"""Build libftl_b200.so in-tree with nvcc for sm_100a.

`python -m fault_tolerant_llm_training_b200.build` (or `__graft_entry__.build()`).  nvcc cross-compiles
without a GPU; the resulting .so is git-ignored but travels to the GPU box with the repo snapshot.
Objects are rebuilt only when their source (or a common header) is newer.
"""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
CSRC = os.path.join(ROOT, "csrc")
OUT_DIR = os.path.join(PKG_DIR, "_native")
LIB_PATH = os.path.join(OUT_DIR, "libftl_b200.so")
BUILD_DIR = os.path.join(ROOT, "build", "obj")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ARCH + [
    "-lineinfo", "-O3", "-std=c++17", "--expt-relaxed-constexpr",
    "-Xcompiler", "-fPIC,-fvisibility=hidden", "-I", CSRC, "-I", os.path.join(ROOT, "include"),
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _newest_header() -> float:
    hs = glob.glob(os.path.join(CSRC, "**", "*.cuh"), recursive=True) + glob.glob(os.path.join(ROOT, "include", "*.h"))
    return max((os.path.getmtime(h) for h in hs), default=0.0)


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed: {' '.join(cmd)}\n{r.stdout}\n{r.stderr}")
    return r.stderr


def build(verbose: bool = False, force: bool = False) -> str:
    nvcc = find_nvcc()
    os.makedirs(OUT_DIR, exist_ok=True)
    os.makedirs(BUILD_DIR, exist_ok=True)
    sources = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    hdr_time = _newest_header()
    jobs, objs = [], []
    for src in sources:
        obj = os.path.join(BUILD_DIR, os.path.basename(src) + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_time):
            jobs.append([nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj])
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for out in ex.map(_run, jobs):
            if verbose and out:
                print(out, file=sys.stderr)
    if jobs or not os.path.exists(LIB_PATH):
        tmp = LIB_PATH + ".tmp"
        _run([nvcc, "-shared"] + ARCH + ["-o", tmp] + objs)
        os.replace(tmp, LIB_PATH)
    return LIB_PATH


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
