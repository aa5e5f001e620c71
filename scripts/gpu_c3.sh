#!/bin/bash
# C3 (non-regular patterns): parity tests of the backtracking path, bench line and launch list. Usage: gpu_c3.sh <tag>
TAG=${1:-x}
mkdir -p gpurun_out; export LP_NO_REBUILD=1
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "c3 or vm or span or backtrack or random or c1 or deep" > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
for nb in 2 4 12; do
  LP_VM_BLOCKS=$nb timeout 300 python bench.py --config c3 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_${TAG}_b$nb.json 2> gpurun_out/bench_${TAG}_b$nb.err; echo "bench b$nb rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${TAG}_b$nb.json").read().strip().splitlines()[-1])
    print("b$nb", d["ms_per_step"], d["phase_ms"], d["parity"]["ok"])
except Exception as e: print("b$nb failed", e)
PY
done
bash scripts/gpu_launches.sh $TAG --config c3
