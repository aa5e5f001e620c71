#!/bin/bash
# A/B of one environment switch on one B200: parity subset once, then a bench line per value (C2 and C3).
# Usage: gpu_ab_env.sh <tag> <ENV_VAR> "<values>"
TAG=${1:-x}; VAR=$2; VALS=$3
mkdir -p gpurun_out; export LP_NO_REBUILD=1
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "c1 or c2 or c3 or random or medium" > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_$TAG.log
for v in $VALS; do
  for cfg in c2; do
    env $VAR=$v timeout 300 python bench.py --config $cfg --steps 20 --warmup 5 --no-cpu-baseline --no-e2e > gpurun_out/bench_${TAG}_${v}_$cfg.json 2> gpurun_out/bench_${TAG}_${v}_$cfg.err; echo "bench $VAR=$v $cfg rc=$?"
    python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${TAG}_${v}_$cfg.json").read().strip().splitlines()[-1])
    print("$VAR=$v $cfg", round(d["ms_per_step"],4), {k:round(x,3) for k,x in d["phase_ms"].items()}, d["parity"]["ok"])
except Exception as e: print("failed", e)
PY
  done
done
