#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 1200 python -m pytest tests -x -q -p no:cacheprovider --timeout 900 -m gpu > gpurun_out/t_gpu.log 2>&1; echo "gpu tests rc=$?"
tail -n 2 gpurun_out/t_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1
timeout 600 python bench.py --no-cpu-baseline 2>/dev/null | cut -c1-200
