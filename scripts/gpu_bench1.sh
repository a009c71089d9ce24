#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"
tail -n 2 gpurun_out/bench.err | cut -c1-200
grep -h '^{' gpurun_out/bench.log | cut -c1-300
