#!/bin/bash
# bench lines of the other BASELINE.json configs on one B200 (C1 latency, C3 non-regular patterns, C4 scoring-heavy). Usage: gpu_configs.sh <tag>
TAG=${1:-x}
mkdir -p gpurun_out; export LP_NO_REBUILD=1
for cfg in c1 c3 c4; do
  timeout 600 python bench.py --config $cfg --steps 20 --warmup 5 > gpurun_out/bench_${TAG}_$cfg.json 2> gpurun_out/bench_${TAG}_$cfg.err; echo "bench $cfg rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${TAG}_$cfg.json").read().strip().splitlines()[-1])
    print("$cfg", d["ms_per_step"], d["phase_ms"], d["e2e"]["value"], d["parity"], d["cpu_baseline"]["value"])
except Exception as e: print("$cfg failed", e)
PY
done
