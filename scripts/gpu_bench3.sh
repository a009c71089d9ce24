#!/bin/bash
mkdir -p gpurun_out
run() {
  local t0=$(date +%s.%N)
  timeout 600 python bench.py --no-cpu-baseline 2> gpurun_out/err.txt | python -c 'import sys,json; d=json.loads(sys.stdin.readline()); print(sys.argv[1], d["steps"], "steps", round(d["ms_per_step"],3), "ms/step", "e2e", round(d["e2e"]["ms_per_step"],3), (d["clocks"] or {}).get("samples"), d["gpu_launches"])' "$1"
  echo "wall $(echo "$(date +%s.%N) - $t0" | bc) s"
}
for i in 1 2 3 4; do run "default:"; done
tail -n 3 gpurun_out/err.txt
