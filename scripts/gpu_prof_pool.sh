#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 300 python scripts/kernel_bench.py --only relu_ --prof > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"relu_|bias_fold" -c 40 -o gpurun_out/prof_pool \
   python scripts/kernel_bench.py --only relu_ --prof > gpurun_out/ncu_pool.log 2>&1; echo "ncu rc=$?"
