#!/bin/bash
mkdir -p gpurun_out
export PYTHONUNBUFFERED=1
timeout 600 python -c "import __graft_entry__ as g; g.build(); g.smoke(); print('SMOKE OK')" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
tail -n 6 gpurun_out/smoke.log | cut -c1-300
