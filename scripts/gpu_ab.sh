#!/bin/bash
# A/B of the scan variants on one B200: parity tests first, then the bench line of each variant.
mkdir -p gpurun_out; export LP_NO_REBUILD=1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
tail -3 gpurun_out/smoke.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
for v in 0 1 2; do
  LP_FSCAN_VARIANT=$v timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_fused_v$v.json 2> gpurun_out/bench_fused_v$v.err; echo "bench v$v rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_fused_v$v.json").read().strip().splitlines()[-1])
    print("v$v", d["ms_per_step"], d["phase_ms"], d["roofline"]["frac"], d["candidates"], d["e2e"]["value"])
except Exception as e: print("v$v failed", e)
PY
done
LP_SCAN_MODE=legacy timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_legacy.json 2> gpurun_out/bench_legacy.err; echo "bench legacy rc=$?"
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_legacy.json").read().strip().splitlines()[-1])
    print("legacy", d["ms_per_step"], d["phase_ms"], d["roofline"]["frac"], d["candidates"], d["e2e"]["value"])
except Exception as e: print("legacy failed", e)
PY
