#!/bin/bash
# per-kernel device times of a few steps (cold-cache, serialised: compare shares). Usage: gpu_launches.sh <tag> [bench args]
TAG=${1:-x}; shift
mkdir -p gpurun_out; export LP_NO_REBUILD=1
timeout 300 python bench.py --steps 20 --warmup 5 "$@" > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_$TAG.json").read().strip().splitlines()[-1])
print("$TAG", d["ms_per_step"], d["phase_ms"], d["roofline"]["frac"], d["e2e"]["value"], d.get("parity"))
PY
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-parity "$@" > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --no-parity "$@" > gpurun_out/ncu_$TAG.log 2>&1; echo "ncu rc=$?"
