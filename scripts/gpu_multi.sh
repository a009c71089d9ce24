#!/bin/bash
# usage: bash scripts/gpu_multi.sh N   (run under gpurun --gpus N)
N=${1:-2}
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi -L > gpurun_out/gpus.txt 2>&1
timeout 900 python -m pytest -q -p no:cacheprovider --timeout 600 tests/test_gpu_ddp.py -m gpu > gpurun_out/t_ddp.log 2>&1; echo "ddp tests rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.log 2>&1; echo "bench N=$N rc=$?"
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n1.log 2>&1; echo "bench N=1 rc=$?"
tail -n 15 gpurun_out/t_ddp.log
grep -h '^{' gpurun_out/bench_n$N.log | cut -c1-400
grep -h '^{' gpurun_out/bench_n1.log | cut -c1-200
tail -n 5 gpurun_out/bench_n$N.log | cut -c1-300
