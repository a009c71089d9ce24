"""Per-kernel shares of the bench's timed steps from an ncu launch list (gpu__time_duration.sum, CSV):
    python scripts/launch_shares.py gpurun_out/launches.csv > profiles/rNN_bench_step_shares.csv
Times are per launch, cold-cache and serialised by ncu: shares are meaningful, absolutes are not."""
import csv
import re
import sys
from collections import OrderedDict

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
agg = OrderedDict()
for r in rows:
    name = re.sub(r"\(.*", "", r[4]).replace("void ", "").strip()
    name = re.sub(r"b200::(<unnamed>::)?", "", name)
    us = float(r[-1]) / (1000.0 if r[-2] in ("ns", "nsecond") else 1.0)
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += us
tot = sum(v[1] for v in agg.values())
own = sum(v[1] for k, v in agg.items() if not k.startswith("at::") and "nccl" not in k.lower())
print("kernel,launches,total_us,share")
for k, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"\"{k}\",{n},{us:.1f},{us / tot:.4f}")
print(f"\"TOTAL\",{sum(v[0] for v in agg.values())},{tot:.1f},1.0")
print(f"\"own kernels (not at:: / nccl)\",,{own:.1f},{own / tot:.4f}")
