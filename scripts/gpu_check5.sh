#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
T="timeout 900 python -m pytest -q -p no:cacheprovider --timeout 600"
$T tests/test_gpu_conv.py tests/test_gpu_step.py -m gpu > gpurun_out/t_conv_step.log 2>&1; echo "conv+step rc=$?"
timeout 600 python scripts/kernel_bench.py --only conv2d,linear > gpurun_out/kernel_bench_conv.log 2>&1; echo "kbench rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench.log 2>&1; echo "bench rc=$?"
tail -n 4 gpurun_out/t_conv_step.log
python - <<'PY'
import json
for l in open('gpurun_out/kernel_bench_conv.log'):
    if l.startswith('{'):
        d = json.loads(l); print(f"{d['ms']*1000:8.1f} us  {str(d['TFLOPs'] or ''):>7}  {d['kernel']}")
    else: print(l.rstrip()[:200])
PY
tail -n 1 gpurun_out/bench.log | cut -c1-300
