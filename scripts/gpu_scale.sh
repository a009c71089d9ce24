#!/bin/bash
# usage: bash scripts/gpu_scale.sh N   (under gpurun --gpus N): the multi-GPU DDP tests are skipped, bench at N and at 1
N=${1:-8}
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n$N.log 2>&1; echo "bench N=$N rc=$?"
grep -h '^{' gpurun_out/bench_n$N.log | cut -c1-260
tail -n 3 gpurun_out/bench_n$N.log | cut -c1-300
