"""In-tree build of liblogparser_b200.so for sm_100a (nvcc cross-compiles without a GPU)."""
from __future__ import annotations

import hashlib
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


NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-fmad=false",  # Java is strict IEEE-754 without contraction (JEP 306); scores must not use FMA
              "-Xcompiler", "-fPIC", "-shared"]
LAST_BUILD_MODE = None      # "compiled" | "prebuilt": what build() did in this process (bench.py reports it)


def source_hash(sources=None, flags=None) -> str:
    """Fingerprint of everything that goes into a library: sources, headers and compiler flags. Stored beside the
    .so (<name>.srchash) so that a copy of the tree on another box, whose file times mean nothing, reuses the
    prebuilt library exactly when it was built from these sources."""
    h = hashlib.sha256()
    for f in (sources or SOURCES) + HEADERS:
        p = os.path.join(CSRC, f)
        h.update(f.encode())
        if os.path.exists(p):
            with open(p, "rb") as fh:
                h.update(fh.read())
    h.update(" ".join(flags if flags is not None else NVCC_FLAGS).encode())
    return h.hexdigest()[:16]


def stale(so: str = SO, sources=None, flags=None) -> bool:
    if not os.path.exists(so) or not os.path.exists(so + ".srchash"):
        return True
    with open(so + ".srchash") as fh:
        return fh.read().strip() != source_hash(sources, flags)


def build(force: bool = False, verbose: bool = False) -> str:
    global LAST_BUILD_MODE
    if not force and (not stale() or (os.environ.get("LP_NO_REBUILD") and os.path.exists(SO))):
        LAST_BUILD_MODE = LAST_BUILD_MODE or "prebuilt"
        return SO       # LP_NO_REBUILD: a GPU run uses the library that travelled with the snapshot, whatever the sources say
    cmd = [_nvcc()] + NVCC_FLAGS + ["-o", SO] + [os.path.join(CSRC, s) for s in SOURCES] + ["-ldl"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    with open(SO + ".srchash", "w") as fh:
        fh.write(source_hash() + "\n")
    LAST_BUILD_MODE = "compiled"
    if verbose:
        print(r.stdout + r.stderr)
    return SO


DEBUG_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-shared"]


def build_debug(force: bool = False) -> str:
    """liblogparser_b200_debug.so: lp_debug_* (DebugRegex, plan dump, scan-group walk) for the CPU test tier."""
    if not force and not stale(DEBUG_SO, DEBUG_SOURCES, DEBUG_FLAGS):
        return DEBUG_SO
    cxx = shutil.which("g++") or "g++"
    cmd = [cxx] + DEBUG_FLAGS + ["-o", DEBUG_SO] + [os.path.join(CSRC, s) for s in DEBUG_SOURCES]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("g++ failed:\n" + r.stdout + r.stderr)
    with open(DEBUG_SO + ".srchash", "w") as fh:
        fh.write(source_hash(DEBUG_SOURCES, DEBUG_FLAGS) + "\n")
    return DEBUG_SO


if __name__ == "__main__":
    print(build(force=True, verbose=True))
    print(build_debug(force=True))
