"""Builds libyolohead.so (hand-written sm_100a CUDA behind a C ABI) in-tree with nvcc.

    python -m yolo_v1_b200.build [--force] [--verbose]

The .so lands next to this file so it travels with the repo snapshot to the GPU
box; nothing is JIT-compiled at import time.
"""
from __future__ import annotations

import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libyolohead.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",            # one rounding per operator, like the reference's Python / ATen math
    "-prec-div=true", "-prec-sqrt=true",
    "-Xcompiler", "-fPIC", "-shared",
    "--use-local-env", "-ccbin", "/usr/bin/g++",
]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def is_stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.join(HERE, "..", "include", "yolohead.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False, out: str | None = None) -> str:
    extra = os.environ.get("YH_NVCC_EXTRA", "").split()  # experiments only, e.g. -DYH_SORT_LOOK=2
    if out is not None:                                  # an experiment build next to the product library
        return _compile(out, extra, verbose)
    flags_file = LIB + ".flags"                          # a change of experiment flags is a rebuild too
    last = open(flags_file).read() if os.path.exists(flags_file) else ""
    if not force and not is_stale() and last == " ".join(extra):
        return LIB
    _compile(LIB, extra, verbose)
    with open(flags_file, "w") as f:
        f.write(" ".join(extra))
    return LIB


def _compile(target: str, extra, verbose: bool) -> str:
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + list(extra) + (["-Xptxas", "-v"] if verbose else []) + ["-o", target] + sources()
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed building {os.path.basename(target)}")
    return target


if __name__ == "__main__":
    _out = sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else None
    print(build_library(force="--force" in sys.argv, verbose="--verbose" in sys.argv, out=_out))
