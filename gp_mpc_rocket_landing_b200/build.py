"""Build libgpmpc_b200.so (the C-ABI of include/gpmpc_b200.h) in-tree with nvcc for sm_100a.

    python -m gp_mpc_rocket_landing_b200.build [--force]

The library is built next to this file so that it travels to the GPU box with the repo snapshot.
nvcc cross-compiles without a GPU; nothing here needs torch.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libgpmpc_b200.so")
SOURCES = ["gram.cu", "linalg.cu", "admm.cu", "admm_stage.cu", "admm3.cu", "mpc.cu", "mpc6.cu", "sparse.cu", "lml_grad.cu"]
# element-wise model arithmetic must round like the numpy reference (no FMA contraction); the hot inner products in
# these files are DMMA inline asm / explicit fma() and are unaffected
EXTRA_FLAGS = {"admm3.cu": (["-DADMM3_TRACE"] if os.environ.get("ADMM3_TRACE") else []),
               "gram.cu": ["-fmad=false"], "mpc.cu": ["-fmad=false"], "mpc6.cu": ["-fmad=false"], "sparse.cu": ["-fmad=false"]}
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-I" + os.path.join(ROOT, "include"), "-I" + CSRC,
]


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found: the CUDA extension cannot be built (no CPU fallback exists)")


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    nvcc = _nvcc()
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(ROOT, "include", "gpmpc_b200.h"))
    jobs = []
    objs = []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ, s.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [src] + headers):
            jobs.append([nvcc] + NVCC_FLAGS + EXTRA_FLAGS.get(s, []) + (["-Xptxas", "-v"] if verbose else [])
                        + ["-c", src, "-o", obj])

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        return r.stderr

    with ThreadPoolExecutor(max_workers=4) as ex:
        for out in ex.map(run, jobs):
            if verbose and out:
                print(out)
    if jobs or force or _stale(LIB, objs):
        run([nvcc, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
