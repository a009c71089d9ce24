#!/bin/bash
# One gpurun call: GPU parity tests (strict-fp32 paths first, tcgen05 paths in their own processes so a
# trap cannot poison the others), smoke, a short bench, and the ncu launch list of the same bench command.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
T="timeout 900 python -m pytest -q -p no:cacheprovider --timeout 600"
$T tests/test_gpu_kernels.py -m gpu -k "not tf32" > gpurun_out/t_kernels_fp32.log 2>&1; echo "kernels_fp32 rc=$?"
$T tests/test_gpu_kernels.py -m gpu -k "tf32" > gpurun_out/t_kernels_tf32.log 2>&1; echo "kernels_tf32 rc=$?"
$T tests/test_gpu_step.py -m gpu -k "not tf32" > gpurun_out/t_step_fp32.log 2>&1; echo "step_fp32 rc=$?"
$T tests/test_gpu_step.py -m gpu -k "tf32" > gpurun_out/t_step_tf32.log 2>&1; echo "step_tf32 rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 --mode fp32 --no-cpu-baseline > gpurun_out/bench_fp32.log 2>&1; echo "bench_fp32 rc=$?"
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?"
tail -n 5 gpurun_out/t_*.log gpurun_out/smoke.log
tail -n 2 gpurun_out/bench.log | cut -c1-1500
