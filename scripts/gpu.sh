#!/bin/bash
# The gpurun command lines of this repo, in one place:  gpurun [--gpus N] --timeout T -- 'bash scripts/gpu.sh <what> [args]'
# Everything lands in gpurun_out/ (scratch); summaries worth keeping are copied to profiles/ by hand.
#   tests [pytest args]  GPU parity tests on one GPU (not multigpu), then smoke()
#   multi N              multigpu tests (2 ranks) + bench.py at N ranks
#   bench [args]         bench.py with the given arguments
#   launches             ncu launch list of a 3-step bench (after the same command ran clean without ncu)
#   ncu REGEX N [args]   ncu --set full of the kernels matching REGEX (N launches) from scripts/kernel_bench.py
#   kbench [args]        scripts/kernel_bench.py
#   sweep N VAR v1 v2..  bench.py at N ranks once per value of the environment variable VAR (timed regions only)
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
what=$1; shift
T="timeout 1500 python -m pytest -q -p no:cacheprovider --timeout 900"
case "$what" in
  tests)
    nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/gpu.txt 2>&1
    $T tests -m "gpu and not multigpu" "$@" > gpurun_out/t_gpu.log 2>&1; echo "tests rc=$?"
    tail -n 40 gpurun_out/t_gpu.log
    timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
    ;;
  multi)
    n=${1:-2}
    $T tests -m "multigpu" > gpurun_out/t_multi.log 2>&1; echo "multi tests rc=$?"; tail -n 25 gpurun_out/t_multi.log
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $n --steps 50 --warmup 5 > gpurun_out/bench_n$n.log 2>gpurun_out/bench_n$n.err; echo "bench n=$n rc=$?"
    grep -h '^{' gpurun_out/bench_n$n.log | cut -c1-4000; tail -n 8 gpurun_out/bench_n$n.err
    ;;
  bench)
    timeout 900 python bench.py "$@" > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"
    grep -h '^{' gpurun_out/bench.log | cut -c1-6000; tail -n 8 gpurun_out/bench.err
    ;;
  launches)
    timeout 300 python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/bench_short.log 2>&1 && \
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv \
       python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/ncu_bench.log 2>&1; echo "ncu list rc=$?"
    ;;
  ncu)
    regex=$1; count=${2:-8}; shift 2
    timeout 300 python scripts/kernel_bench.py --prof "$@" > gpurun_out/prof_plain.log 2>&1 && \
    timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$regex -c $count -o gpurun_out/prof_$regex -f \
       python scripts/kernel_bench.py --prof "$@" > gpurun_out/ncu_$regex.log 2>&1; echo "ncu rc=$?"
    ;;
  kbench)
    timeout 900 python scripts/kernel_bench.py "$@" > gpurun_out/kernel_bench.log 2>&1; echo "kbench rc=$?"; tail -n 60 gpurun_out/kernel_bench.log
    ;;
  sweep)
    # sweep N VAR v1 v2 ... : bench.py at N ranks (30 steps, timed regions only) for each value of environment variable VAR
    n=$1; var=$2; shift 2
    for v in "$@"; do
      if [ "$n" = "1" ]; then
        env $var=$v timeout 600 python bench.py --steps 100 --warmup 5 --no-extras > gpurun_out/sweep_${var}_$v.log 2>gpurun_out/sweep_${var}_$v.err
      else
        env $var=$v timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 \
          bench.py --gpus $n --steps 30 --warmup 5 --no-extras $BENCH_ARGS > gpurun_out/sweep_${var}_$v.log 2>gpurun_out/sweep_${var}_$v.err
      fi
      python - "$var=$v" gpurun_out/sweep_${var}_$v.log <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[2]) if l.startswith("{")][-1])
    n = d.get("nccl") or {}
    print(sys.argv[1], "ms/step %.3f" % d["ms_per_step"], "e2e %.3f" % d["e2e"]["ms_per_step"], "coll_ms %s busbw %s exposed %s dry %s" % (
        n.get("collective_ms_per_step"), n.get("busbw_GBps"), n.get("exposed_comm_ms"), n.get("dry_step_ms")))
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
    done
    ;;
  *) echo "unknown: $what"; exit 2;;
esac
