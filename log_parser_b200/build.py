"""In-tree build of liblogparser_b200.so for sm_100a (nvcc cross-compiles without a GPU)."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "liblogparser_b200.so")
SOURCES = ["engine.cu", "noalloc.cu", "plan.cpp", "rx_parse.cpp", "rx_compile.cpp", "rx_literal.cpp", "vm.cpp"]
# test-only introspection of the host compiler: its own library, host code only (g++), never loaded by the product path
DEBUG_SO = os.path.join(HERE, "liblogparser_b200_debug.so")
DEBUG_SOURCES = ["debug_api.cpp", "plan.cpp", "rx_parse.cpp", "rx_compile.cpp", "rx_literal.cpp", "vm.cpp"]
HEADERS = ["kernels.cuh", "k_scan.cuh", "k_fscan.cuh", "k_verify.cuh", "k_score.cuh", "k_sort.cuh", "plan.h", "rx.h", "vm.h", "vm_exec.h", os.path.join("..", "..", "include", "logparser_b200.h")]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: liblogparser_b200.so cannot be built (there is no CPU fallback)")


def stale(so: str = SO, sources=None) -> bool:
    if not os.path.exists(so):
        return True
    t = os.path.getmtime(so)
    for f in (sources or SOURCES) + HEADERS:
        p = os.path.join(CSRC, f)
        if os.path.exists(p) and os.path.getmtime(p) > t:
            return True
    return False


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and (not stale() or (os.environ.get("LP_NO_REBUILD") and os.path.exists(SO))):
        return SO       # LP_NO_REBUILD: a GPU run uses the library that travelled with the snapshot, whatever the source mtimes say
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


def build_debug(force: bool = False) -> str:
    """liblogparser_b200_debug.so: lp_debug_* (DebugRegex, plan dump, scan-group walk) for the CPU test tier."""
    if not force and not stale(DEBUG_SO, DEBUG_SOURCES):
        return DEBUG_SO
    cxx = shutil.which("g++") or "g++"
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-o", DEBUG_SO] + [os.path.join(CSRC, s) for s in DEBUG_SOURCES]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("g++ failed:\n" + r.stdout + r.stderr)
    return DEBUG_SO


if __name__ == "__main__":
    print(build(force=True, verbose=True))
    print(build_debug(force=True))
