"""Builds the C-ABI CUDA library in-tree: ``ivit_b200/lib/libivit_b200.so`` (sm_100a only).

nvcc cross-compiles without a GPU; the library links only the (static) CUDA runtime, so it
also *loads* on a CPU-only box (symbol checks), and fetches ``cuTensorMapEncodeTiled`` from the
driver at run time.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libivit_b200.so")
SOURCES = ["rowwise.cu", "rowwise16.cu", "swin.cu", "gemm_i8.cu", "attention.cu"]
HEADERS = ["ptx.cuh", "intmath.cuh", "launch.cuh"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--fmad=false",            # the exact-replay arithmetic must never be contracted into FMAs
    "-I" + os.path.join(ROOT, "include"), "-I" + CSRC,
] + os.environ.get("IVIT_NVCC_EXTRA", "").split()          # e.g. -DIVIT_ATTN_TIMING for the per-phase clock probe


def _newest(paths):
    return max(os.path.getmtime(p) for p in paths)


def needs_build() -> bool:
    if not os.path.isfile(LIB_PATH):
        return True
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.join(ROOT, "include", "ivit_b200.h"),
                                                                  os.path.abspath(__file__)]
    return _newest(deps) > os.path.getmtime(LIB_PATH)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    os.makedirs(LIB_DIR, exist_ok=True)
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(LIB_DIR, src.replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n{out}")
        if verbose and out:
            print(out)
    link = [nvcc, "-shared", "-o", LIB_PATH, *objs, "-gencode", "arch=compute_100a,code=sm_100a",
            "-cudart", "static", "-Xcompiler", "-fPIC"]
    r = subprocess.run(link, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}")
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
