#!/bin/bash
mkdir -p gpurun_out
timeout 600 python scripts/kernel_bench.py --only "conv2d_dgrad conv" > gpurun_out/kb_dgrad.log 2>&1; echo "rc=$?"
timeout 600 python scripts/kernel_bench.py --only "conv2d_fwd conv" >> gpurun_out/kb_dgrad.log 2>&1; echo "rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/kb_dgrad.log'):
    if l.startswith('{'):
        d = json.loads(l); print(f"{d['ms']*1000:8.1f} us  {str(d.get('TFLOPs') and round(d['TFLOPs'])):>6} TF  {d['kernel']}")
PY
