#!/bin/bash
# Round-end measurement pass on one B200: full GPU test suite, both bench arms, the ncu launch list of the bench
# command and one `--set full` capture of the heaviest kernels. Everything lands in gpurun_out/.
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 1500 python -m pytest tests -q -p no:cacheprovider --timeout 900 -m "gpu and not multigpu" > gpurun_out/t_gpu.log 2>&1; echo "tests rc=$?"
tail -n 3 gpurun_out/t_gpu.log
timeout 600 python bench.py > gpurun_out/bench.log 2>&1; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "ref rc=$?"
timeout 600 python scripts/kernel_bench.py > gpurun_out/kernel_bench.log 2>&1; echo "kbench rc=$?"
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1; echo "ncu list rc=$?"
timeout 300 python scripts/kernel_bench.py --only conv2d --prof > gpurun_out/prof_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:conv_ -c 26 -o gpurun_out/prof_conv \
   python scripts/kernel_bench.py --only conv2d --prof > gpurun_out/ncu_conv.log 2>&1; echo "ncu conv rc=$?"
timeout 300 python scripts/kernel_bench.py --only adam --prof > gpurun_out/prof_plain2.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:adam -c 2 -o gpurun_out/prof_adam \
   python scripts/kernel_bench.py --only adam --prof > gpurun_out/ncu_adam.log 2>&1; echo "ncu adam rc=$?"
grep -h '^{' gpurun_out/bench.log | cut -c1-600
grep -h '^{' gpurun_out/bench_ref.log | cut -c1-600
