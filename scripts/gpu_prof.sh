#!/bin/bash
# bench line + ncu capture of the scan kernel (one GPU). Usage: gpu_prof.sh <tag> [kernel regex]
TAG=${1:-x}; KRE=${2:-k_fscan}
mkdir -p gpurun_out; export LP_NO_REBUILD=1
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "split or c1 or batch or range or piecewise" > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_$TAG.json").read().strip().splitlines()[-1])
print("$TAG", d["ms_per_step"], d["phase_ms"], d["roofline"]["frac"], d["candidates"], d["e2e"]["value"])
PY
timeout 300 python bench.py --steps 2 --warmup 1 > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KRE -s 2 -c 1 -f -o gpurun_out/prof_$TAG python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_$TAG.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_$TAG.log
