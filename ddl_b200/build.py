"""In-tree build of the native library for sm_100a (``python -m ddl_b200.build``).

nvcc is invoked directly -- no torch extension machinery, no JIT cache -- so the resulting
``ddl_b200/lib/libddl_b200.so`` travels with the source tree to the GPU box.  cudart is linked
statically and driver entry points are resolved at run time (cudaGetDriverEntryPoint), so the
library also loads on a machine without a GPU driver.  The reference has no build step beyond
Poetry (reference pyproject.toml:1-30).
"""
from __future__ import annotations

import glob
import hashlib
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "csrc")
LIB_DIR = os.path.join(ROOT, "ddl_b200", "lib")
LIB = os.path.join(LIB_DIR, "libddl_b200.so")
STAMP = os.path.join(LIB_DIR, "libddl_b200.stamp")

NVCC_FLAGS = [
    "-gencode",
    "arch=compute_100a,code=[sm_100a,compute_100a]",
    "-lineinfo",
    "-O3",
    "-std=c++17",
    "-shared",
    "-Xcompiler",
    "-fPIC",
    "-Xcompiler",
    "-pthread",
]


def sources() -> list[str]:
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cpp")))


def _digest() -> str:
    h = hashlib.sha256()
    files = sources() + sorted(glob.glob(os.path.join(CSRC, "*.h")) + glob.glob(os.path.join(CSRC, "*.cuh")))
    files += sorted(glob.glob(os.path.join(ROOT, "include", "*.h")))
    for f in files:
        h.update(f.encode())
        with open(f, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def nvcc_path() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; the native library cannot be built")


def build_native(force: bool = False, verbose: bool = True, extra_flags: list[str] | None = None) -> str:
    """Compile every CUDA/C++ source under csrc/ into one shared object.  Returns its path."""
    os.makedirs(LIB_DIR, exist_ok=True)
    digest = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP):
        with open(STAMP) as fh:
            if fh.read().strip() == digest:
                return LIB
    cmd = [nvcc_path(), *NVCC_FLAGS, *(extra_flags or []), "-I", os.path.join(ROOT, "include"), "-o", LIB, *sources()]
    if verbose:
        print("[ddl_b200.build]", " ".join(cmd), flush=True)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed building libddl_b200.so")
    if verbose and (res.stdout or res.stderr):
        print(res.stdout + res.stderr)
    with open(STAMP, "w") as fh:
        fh.write(digest)
    return LIB


if __name__ == "__main__":
    build_native(force="--force" in sys.argv, extra_flags=["-Xptxas", "-v"] if "--verbose" in sys.argv else None)
    print(LIB)
