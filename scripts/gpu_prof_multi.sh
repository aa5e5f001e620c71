#!/bin/bash
# one `ncu --set full` capture of several kernels of one bench step. Usage: gpu_prof_multi.sh <tag> <kernel regex> <count> [bench args]
TAG=${1:-x}; KRE=${2:-k_}; CNT=${3:-3}; shift; shift; shift
mkdir -p gpurun_out; export LP_NO_REBUILD=1
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity "$@" > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$KRE -s $((CNT*4)) -c $CNT -f -o gpurun_out/prof_$TAG python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-parity "$@" > gpurun_out/ncu_$TAG.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_$TAG.log
