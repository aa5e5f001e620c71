#!/bin/bash
# full GPU parity suite, bench line, per-kernel launch list (one GPU). Usage: gpu_full2.sh <tag>
TAG=${1:-x}
mkdir -p gpurun_out; export LP_NO_REBUILD=1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_$TAG.log
bash scripts/gpu_launches.sh $TAG
