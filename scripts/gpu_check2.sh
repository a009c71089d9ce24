#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
T="timeout 900 python -m pytest -q -p no:cacheprovider --timeout 600"
$T tests/test_gpu_kernels.py -m gpu -k "not tf32" > gpurun_out/t_kernels_fp32.log 2>&1; echo "kernels_fp32 rc=$?"
$T tests/test_gpu_kernels.py -m gpu -k "tf32" > gpurun_out/t_kernels_tf32.log 2>&1; echo "kernels_tf32 rc=$?"
$T tests/test_gpu_step.py -m gpu -s > gpurun_out/t_step.log 2>&1; echo "step rc=$?"
timeout 600 python scripts/kernel_bench.py > gpurun_out/kernel_bench.log 2>&1; echo "kbench rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"
timeout 300 python scripts/kernel_bench.py --only gemm --prof > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tf32_gemm -c 18 -o gpurun_out/prof_gemm \
   python scripts/kernel_bench.py --only gemm --prof > gpurun_out/ncu_gemm.log 2>&1; echo "ncu rc=$?"
tail -n 4 gpurun_out/t_*.log
cat gpurun_out/kernel_bench.log
tail -n 1 gpurun_out/bench.log | cut -c1-400
