#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -m pytest -q -p no:cacheprovider --timeout 300 tests/test_zz_no_alloc.py -m gpu > gpurun_out/t_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"
timeout 1200 python -m pytest -q -p no:cacheprovider --timeout 600 tests -m gpu -x > gpurun_out/t_all_gpu.log 2>&1; echo "all gpu tests rc=$?"
tail -n 3 gpurun_out/t_noalloc_alone.log; tail -n 6 gpurun_out/t_all_gpu.log
