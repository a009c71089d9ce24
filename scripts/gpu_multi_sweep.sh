#!/bin/bash
# usage: bash scripts/gpu_multi_sweep.sh N   (under gpurun --gpus N): bench at N GPUs for several NCCL channel caps
N=${1:-2}
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
for CH in default 6 12 16; do
  if [ "$CH" = "default" ]; then unset NCCL_MAX_NCHANNELS; else export NCCL_MAX_NCHANNELS=$CH; fi
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n${N}_ch$CH.log 2>&1
  echo "channels=$CH rc=$? $(grep -h '^{' gpurun_out/bench_n${N}_ch$CH.log | python -c 'import sys,json; d=json.loads(sys.stdin.readline()); print(round(d["value"]), "samples/s", round(d["ms_per_step"],3), "ms/step")' 2>/dev/null)"
done
