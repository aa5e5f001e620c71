#!/bin/bash
# full GPU parity suite, then the C1 bench line and launch list (one GPU). Usage: gpu_c1.sh <tag>
TAG=${1:-x}
mkdir -p gpurun_out; export LP_NO_REBUILD=1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
bash scripts/gpu_launches.sh $TAG --config c1
