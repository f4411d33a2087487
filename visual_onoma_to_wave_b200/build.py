"""Build libvo2w.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo).

Every translation unit is compiled to an object on its own (in parallel, only when it or a header is newer than
its object), then linked: `nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -c x.cu`, `nvcc -shared`.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_build" + ("_x" if os.environ.get("VO2W_LIB_OUT") else ""))
LIB = os.environ.get("VO2W_LIB_OUT") or os.path.join(HERE, "libvo2w.so")   # VO2W_LIB_OUT + VO2W_NVCC_EXTRA: experiment builds
SOURCES = ["errors.cu", "lr.cu", "conv_simt.cu", "conv_tf32.cu", "conv_umma.cu", "pack.cu", "pair_umma.cu", "generator.cu",
           "rows.cu", "postnet.cu", "decoder.cu", "variance.cu", "stream_umma.cu"]
HEADERS = ["common.cuh", "conv_simt.cuh", "conv_tf32.cuh", "conv_umma.cuh", "pack.cuh", "pair_umma.cuh", "rows.cuh",
           "stream_umma.cuh", "umma_cg.cuh", "../../include/vo2w.h"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]


def _mtime(p: str) -> float:
    return os.path.getmtime(p) if os.path.exists(p) else 0.0


def _stale() -> bool:
    t = _mtime(LIB)
    return t == 0.0 or any(_mtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("VO2W_NVCC_EXTRA", "").split()         # e.g. -DVO2W_EXPERIMENTS for what-if timing builds
    os.makedirs(OBJ, exist_ok=True)
    hdr_t = max(_mtime(os.path.join(CSRC, h)) for h in HEADERS)
    jobs = []
    for s in SOURCES:
        src, obj = os.path.join(CSRC, s), os.path.join(OBJ, s[:-3] + ".o")
        if force or extra or _mtime(obj) < max(_mtime(src), hdr_t):
            cmd = [nvcc, *NVCC_FLAGS, *extra, "-c", src, "-o", obj]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)

    def run(cmd):
        if verbose:
            print(" ".join(cmd))
        return subprocess.run(cmd, capture_output=True, text=True)

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as pool:
        results = list(pool.map(run, jobs))
    for cmd, r in zip(jobs, results):
        if verbose or r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
        if r.returncode != 0:
            raise subprocess.CalledProcessError(r.returncode, cmd)
    objs = [os.path.join(OBJ, s[:-3] + ".o") for s in SOURCES]
    subprocess.run([nvcc, *NVCC_FLAGS, "-shared", "-o", LIB, *objs], check=True)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
