#!/bin/bash
# usage: bash scripts/gpu_scale_sweep.sh N  (under gpurun --gpus N): bench at N GPUs for several NCCL channel caps
N=${1:-8}
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
for CH in 8 16 24 32; do
  export NCCL_MAX_NCHANNELS=$CH
  unset NCCL_DEBUG; unset NCCL_DEBUG_SUBSYS
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n${N}_ch$CH.log 2>&1
  echo "channels=$CH rc=$? $(grep -h '^{' gpurun_out/bench_n${N}_ch$CH.log | python -c 'import sys,json; d=json.loads(sys.stdin.readline()); print(round(d["value"]), "samples/s", round(d["ms_per_step"],3), "ms/step")' 2>/dev/null)"
done
grep -h -i "nvls\|channels\|Algo\|Ring\|Tree" gpurun_out/bench_n${N}_ch16.log | grep -v "^{" | cut -c1-200 | sort | uniq -c | sort -rn | head -30 > gpurun_out/nccl_info_n$N.txt
head -30 gpurun_out/nccl_info_n$N.txt
