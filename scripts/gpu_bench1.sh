#!/bin/bash
mkdir -p gpurun_out
for i in 1; do
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench.log 2>gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/bench.log') if l.startswith('{')][-1])
print(round(d['ms_per_step'],3), 'e2e', round(d['e2e']['ms_per_step'],3), d.get('step_ms_percentiles'))
PY
done
