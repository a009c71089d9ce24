#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
T="timeout 900 python -m pytest -q -p no:cacheprovider --timeout 600"
$T tests/test_gpu_conv.py -m gpu > gpurun_out/t_conv.log 2>&1; echo "conv rc=$?"
$T tests/test_gpu_kernels.py -m gpu > gpurun_out/t_kernels.log 2>&1; echo "kernels rc=$?"
$T tests/test_gpu_step.py -m gpu -s > gpurun_out/t_step.log 2>&1; echo "step rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python scripts/kernel_bench.py > gpurun_out/kernel_bench.log 2>&1; echo "kbench rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?"
tail -n 6 gpurun_out/t_*.log gpurun_out/smoke.log
cat gpurun_out/kernel_bench.log | cut -c1-200
tail -n 1 gpurun_out/bench.log | cut -c1-300
