#!/bin/bash
# CPU only: builds the host compiler (the sources of liblogparser_b200_debug.so: regex parser, automata, literals,
# bytecode, plan) with AddressSanitizer + UndefinedBehaviorSanitizer into gpurun_out/ (scratch) and runs one of the host
# fuzzers against that build. Hunting tool.
# usage: scripts/asan_host_fuzz.sh scripts/fuzz_host_scripts.py 1 3 1000
set -e
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
SO=$PWD/gpurun_out/liblogparser_b200_debug_asan.so
SRC="debug_api.cpp plan.cpp rx_parse.cpp rx_compile.cpp rx_literal.cpp vm.cpp"
if [ ! -f "$SO" ] || [ -n "$(cd log_parser_b200/csrc && find $SRC *.h -newer "$SO" 2>/dev/null)" ]; then
  (cd log_parser_b200/csrc && g++ -O1 -g -fsanitize=address,undefined -fno-omit-frame-pointer -std=c++17 -fPIC -shared -o "$SO" $SRC)
fi
RUNNER=$(mktemp --suffix=.py)
cat > "$RUNNER" <<PY
import runpy, sys
sys.path.insert(0, ".")
from log_parser_b200 import build
build.build_debug = lambda force=False: "$SO"      # the debug library the fuzzers load
sys.argv = sys.argv[1:]
runpy.run_path(sys.argv[0], run_name="__main__")
PY
# python itself is not instrumented: the sanitizer runtime and libstdc++ (for its __cxa_throw hook) come in by preload
LD_PRELOAD="$(g++ -print-file-name=libasan.so) $(g++ -print-file-name=libstdc++.so.6)" ASAN_OPTIONS=detect_leaks=0 \
  UBSAN_OPTIONS=print_stacktrace=1 python "$RUNNER" "$@"
rm -f "$RUNNER"
