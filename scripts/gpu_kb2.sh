#!/bin/bash
mkdir -p gpurun_out
for BN in 256 128; do
  echo "== BN for Cout=256: $BN"
  B200DDP_EXP_BN256=$BN timeout 600 python scripts/kernel_bench.py --only "conv" > gpurun_out/kb_bn.log 2>&1
  python - <<'PY'
import json
for l in open('gpurun_out/kb_bn.log'):
    if l.startswith('{'):
        d = json.loads(l)
        if 'conv4' in d['kernel'] or 'conv5' in d['kernel'] or 'dgrad conv3' in d['kernel']:
            print(f"{d['ms']*1000:8.1f} us  {str(d.get('TFLOPs') and round(d['TFLOPs'])):>6} TF  {d['kernel']}")
PY
done
