#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 300 python scripts/kernel_bench.py --only conv2d --prof > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_ -c 12 -o gpurun_out/prof_conv \
   python scripts/kernel_bench.py --only conv2d --prof > gpurun_out/ncu_conv.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/*.ncu-rep
