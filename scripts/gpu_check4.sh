#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
T="timeout 900 python -m pytest -q -p no:cacheprovider --timeout 600"
$T tests/test_gpu_conv.py -m gpu > gpurun_out/t_conv.log 2>&1; echo "conv rc=$?"
$T tests/test_gpu_kernels.py -m gpu > gpurun_out/t_kernels.log 2>&1; echo "kernels rc=$?"
$T tests/test_gpu_step.py -m gpu -s > gpurun_out/t_step.log 2>&1; echo "step rc=$?"
timeout 300 python scripts/debug_bn.py > gpurun_out/debug_bn.log 2>&1; echo "debug_bn rc=$?"
timeout 600 python scripts/kernel_bench.py > gpurun_out/kernel_bench.log 2>&1; echo "kbench rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.log 2>&1; echo "bench rc=$?"
tail -n 6 gpurun_out/t_*.log
cat gpurun_out/debug_bn.log | tail -40
python - <<'PY'
import json
for l in open('gpurun_out/kernel_bench.log'):
    if l.startswith('{'):
        d = json.loads(l); print(f"{d['ms']*1000:8.1f} us  {d['GBps']:8.1f} GB/s  {d['frac_hbm_peak']:5.2f}  {str(d['TFLOPs'] or ''):>7}  {d['kernel']}")
    else: print(l.rstrip()[:200])
PY
tail -n 1 gpurun_out/bench.log | cut -c1-300
