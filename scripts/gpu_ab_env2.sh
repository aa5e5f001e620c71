#!/bin/bash
# bench lines (C2, device-resident only) for several values of one environment switch. Usage: gpu_ab_env2.sh <tag> <ENV_VAR> "<values>"
TAG=${1:-x}; VAR=$2; VALS=$3
mkdir -p gpurun_out; export LP_NO_REBUILD=1
for v in $VALS; do
  env $VAR=$v timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_${TAG}_${v}.json 2> gpurun_out/bench_${TAG}_${v}.err; echo "bench $VAR=$v rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${TAG}_${v}.json").read().strip().splitlines()[-1])
    print("$VAR=$v", round(d["ms_per_step"],4), {k:round(x,3) for k,x in d["phase_ms"].items()}, d["parity"]["ok"])
except Exception as e: print("failed", e)
PY
done
