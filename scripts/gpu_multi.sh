#!/bin/bash
# N-GPU checks on one box: the NCCL parity test, the C5-shaped bench line (device-resident + host-buffer e2e + parity),
# the reference arm, and the line-range mode on C4. Usage: gpu_multi.sh <N> <tag>
N=${1:-2}; TAG=${2:-m}
mkdir -p gpurun_out; export LP_NO_REBUILD=1
PORT=29517
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "nccl or dist or range" > gpurun_out/pytest_${TAG}.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_${TAG}.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_${TAG}_n$N.json 2> gpurun_out/bench_${TAG}_n$N.err; echo "bench N=$N rc=$?"; tail -c 2500 gpurun_out/bench_${TAG}_n$N.json; tail -5 gpurun_out/bench_${TAG}_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((PORT+1)) bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/ref_${TAG}_n$N.json 2> gpurun_out/ref_${TAG}_n$N.err; echo "reference N=$N rc=$?"; tail -c 600 gpurun_out/ref_${TAG}_n$N.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((PORT+2)) bench.py --gpus $N --steps 10 --warmup 3 --shard lines --config c4 --no-cpu-baseline > gpurun_out/bench_${TAG}_c4lines_n$N.json 2> gpurun_out/bench_${TAG}_c4lines_n$N.err; echo "bench lines N=$N rc=$?"; tail -c 1500 gpurun_out/bench_${TAG}_c4lines_n$N.json; tail -5 gpurun_out/bench_${TAG}_c4lines_n$N.err
