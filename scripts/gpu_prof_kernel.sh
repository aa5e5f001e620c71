#!/bin/bash
# one `ncu --set full` capture of one kernel of a bench run. Usage: gpu_prof_kernel.sh <tag> <kernel regex> [bench args]
TAG=${1:-x}; KRE=${2:-k_fscan}; shift; shift
mkdir -p gpurun_out; export LP_NO_REBUILD=1
export LP_VM_BLOCKS=${LP_VM_BLOCKS:-12}
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-parity "$@" > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KRE -s 2 -c 1 -f -o gpurun_out/prof_$TAG python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-parity "$@" > gpurun_out/ncu_$TAG.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_$TAG.log
