"""Builds ``libspectre_b200.so`` (the C-ABI library of ``include/spectre_b200.h``) in-tree.

``python -m vit_spectre_experiments_b200.build [--force] [--packed-math]``

nvcc cross-compiles for sm_100a without a GPU; the ``.so`` sits next to this file so it
travels with the repo snapshot (it is git-ignored, not gpurun-ignored).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, os.environ.get("SB200_LIB_NAME", "libspectre_b200.so"))   # alternate name: A/B builds
SOURCES = ["fftmix.cu", "fftmix_ring.cu", "fftmix_tc.cu", "dwtmix.cu", "normact.cu", "normact_fast_fwd.cu", "normact_fast_bwd.cu", "permut.cu", "embed.cu",
           "dwtmix_ml.cu", "dwtmix_ml_reg.cu", "haar1d.cu", "noalloc.cu", "capi.cu"]
HEADERS = ["fft_codelets.cuh", "common.cuh", "sm100_ptx.cuh", "normact_common.cuh", "normact_fast.cuh", os.path.join("..", "..", "include", "spectre_b200.h")]
NVCC_FLAGS = [
    "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC",
]


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: cannot build libspectre_b200.so")
    return exe


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, packed_math: bool | None = None, verbose: bool = False) -> str:
    """Compile the three translation units in parallel and link the shared library."""
    if not force and not _stale():
        return LIB
    nvcc = _nvcc()
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    flags = list(NVCC_FLAGS)
    if packed_math is None:
        packed_math = os.environ.get("SB200_PACKED_MATH", "0") == "1"
    flags.append(f"-DSB200_PACKED_MATH={1 if packed_math else 0}")
    flags += os.environ.get("SB200_EXTRA_NVCC_FLAGS", "").split()   # e.g. -DSB200_WALLCLOCK=1 (debug stamps)
    if verbose:
        flags += ["-Xptxas", "-v"]

    def compile_one(src: str) -> str:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = [nvcc, *flags, "-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    tmp = LIB + ".tmp"     # link next to the target, then rename: a reader never sees a half-written library
    cmd = [nvcc, "-shared", "-o", tmp, *objs, "-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    os.replace(tmp, LIB)
    return LIB


if __name__ == "__main__":
    path = build_library(force="--force" in sys.argv, packed_math=True if "--packed-math" in sys.argv else None,
                         verbose="-v" in sys.argv)
    print(path)
