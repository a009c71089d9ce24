#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
T="timeout 1200 python -m pytest -q -p no:cacheprovider --timeout 900 -x"
$T tests -m "gpu and not multigpu" > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"
tail -n 15 gpurun_out/t_gpu.log | cut -c1-220
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench.log 2>&1; echo "bench rc=$?"
tail -n 1 gpurun_out/bench.log | cut -c1-300
