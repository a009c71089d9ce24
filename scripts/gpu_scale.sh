#!/bin/bash
# usage: bash scripts/gpu_scale.sh N   (under gpurun --gpus N): bench at N GPUs (+ which all-reduce algorithm NCCL picked)
N=${1:-8}
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
export NCCL_DEBUG=INFO; export NCCL_DEBUG_SUBSYS=COLL,TUNING
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n$N.log 2>&1; echo "bench N=$N rc=$?"
grep -h '^{' gpurun_out/bench_n$N.log | cut -c1-260
grep -h "Bytes -> Algo" gpurun_out/bench_n$N.log | sed 's/.*AllReduce/AllReduce/' | sort | uniq -c | sort -rn | head -5
