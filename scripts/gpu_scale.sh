#!/bin/bash
# usage: bash scripts/gpu_scale.sh N   (under gpurun --gpus N): bench at N GPUs (+ which all-reduce algorithm NCCL picked)
N=${1:-8}
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
# (NCCL_DEBUG=INFO with the COLL subsystem logs every all-reduce and costs ~0.9 ms per step at 4 GPUs: leave it off for timing)
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n$N.log 2>&1; echo "bench N=$N rc=$?"
grep -h '^{' gpurun_out/bench_n$N.log | cut -c1-260
