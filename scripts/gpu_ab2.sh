#!/bin/bash
# A/B of k_fscan variants on one B200: parity subset, bench line per variant, one ncu capture. Usage: gpu_ab2.sh <tag> "<variants>" <variant to profile>
TAG=${1:-x}; VARS=${2:-"0 3"}; PV=${3:-0}
mkdir -p gpurun_out; export LP_NO_REBUILD=1
for v in $VARS; do
  LP_FSCAN_VARIANT=$v timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q -k "split or c1 or c2 or batch or range or piecewise or noalloc or no_alloc" > gpurun_out/pytest_${TAG}_v$v.log 2>&1; echo "pytest v$v rc=$?"; tail -3 gpurun_out/pytest_${TAG}_v$v.log
  LP_FSCAN_VARIANT=$v timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_${TAG}_v$v.json 2> gpurun_out/bench_${TAG}_v$v.err; echo "bench v$v rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${TAG}_v$v.json").read().strip().splitlines()[-1])
    print("v$v", d["ms_per_step"], d["phase_ms"], d["roofline"]["frac"], d["candidates"], d["e2e"]["value"], d["parity"]["ok"])
except Exception as e: print("v$v failed", e)
PY
done
LP_FSCAN_VARIANT=$PV timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/plain_$TAG.log 2>&1 &&
LP_FSCAN_VARIANT=$PV timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fscan -s 2 -c 1 -f -o gpurun_out/prof_$TAG python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/ncu_$TAG.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_$TAG.log
