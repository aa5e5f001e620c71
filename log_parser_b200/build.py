"""In-tree build of liblogparser_b200.so for sm_100a (nvcc cross-compiles without a GPU)."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "liblogparser_b200.so")
SOURCES = ["engine.cu", "noalloc.cu", "plan.cpp", "rx_parse.cpp", "rx_compile.cpp", "rx_literal.cpp", "vm.cpp", "debug_api.cpp"]
HEADERS = ["kernels.cuh", "k_scan.cuh", "k_verify.cuh", "k_score.cuh", "k_sort.cuh", "plan.h", "rx.h", "vm.h", "vm_exec.h", os.path.join("..", "..", "include", "logparser_b200.h")]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: liblogparser_b200.so cannot be built (there is no CPU fallback)")


def stale() -> bool:
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    for f in SOURCES + HEADERS:
        p = os.path.join(CSRC, f)
        if os.path.exists(p) and os.path.getmtime(p) > t:
            return True
    return False


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        return SO
    cmd = [
        _nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
        "-fmad=false",  # Java is strict IEEE-754 without contraction (JEP 306); scores must not use FMA
        "-Xcompiler", "-fPIC", "-shared", "-o", SO,
    ] + [os.path.join(CSRC, s) for s in SOURCES] + ["-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout + r.stderr)
    return SO


if __name__ == "__main__":
    print(build(force=True, verbose=True))
