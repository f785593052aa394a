"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` log: launches and device time per kernel.
   python bench_tools/launch_census.py gpurun_out/launches_step_fp16.csv [skip_launches] > profiles/ncu/launches_step_fp16_summary.txt"""
import collections
import csv
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 5]
hdr = next(r for r in rows if "Kernel Name" in r)
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
data = [r for r in rows if r is not hdr and len(r) > vi and r[vi].replace(".", "").replace(",", "").isdigit()]
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
data = data[skip:]
agg = collections.OrderedDict()
for r in data:
    name = re.sub(r"\(.*", "", r[ki]).replace("void ", "").replace("pdb::", "")
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += float(r[vi].replace(",", ""))
tot = sum(v[1] for v in agg.values())
print(f"# {len(data)} kernel launches, sum of device durations {tot / 1e6:.3f} ms (ncu --metrics gpu__time_duration.sum, eager launches)")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[1] / 1e6:9.3f} ms  n={v[0]:4d}  {k}")
