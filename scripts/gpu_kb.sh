#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/kernel_bench.py --only relu_ > gpurun_out/kb_relu.log 2>&1; echo "rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/kb_relu.log'):
    if l.startswith('{'):
        d = json.loads(l); print(f"{d['ms']*1000:8.1f} us  {str(d.get('GBps') and round(d['GBps'])):>6} GB/s  {d['kernel']}")
    else: print(l.rstrip()[:200])
PY
