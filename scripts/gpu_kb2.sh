#!/bin/bash
mkdir -p gpurun_out
for ST in 6 3 2; do
  echo "== GEMM stages cap: $ST"
  B200DDP_EXP_GEMM_STAGES=$ST timeout 600 python scripts/kernel_bench.py --only "gemm" > gpurun_out/kb_g.log 2>&1
  python - <<'PY'
import json
for l in open('gpurun_out/kb_g.log'):
    if l.startswith('{'):
        d = json.loads(l); print(f"{d['ms']*1000:8.1f} us  {str(d.get('GBps') and round(d['GBps'])):>6} GB/s  {d['kernel'][:60]}")
PY
done
