"""Condenses an ncu launch list (--csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum)
into one markdown row per kernel: launches, mean time, mean DRAM read / write per launch.
    python tools/launch_summary.py gpurun_out/launches.csv [name-filter-regex] > profiles/launches_rN.md"""
import collections
import csv
import io
import re
import sys

rows = [l for l in open(sys.argv[1]) if not l.startswith("==")]
flt = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
per = collections.OrderedDict()
for d in csv.DictReader(io.StringIO("".join(rows))):
    per.setdefault(d["ID"], {"name": d["Kernel Name"]})[d["Metric Name"]] = float(d["Metric Value"].replace(",", ""))
agg = collections.OrderedDict()
for rec in per.values():
    name = re.sub(r"\(.*", "", rec["name"])
    if flt and not flt.search(name):
        continue
    a = agg.setdefault(name, [0, 0.0, 0.0, 0.0])
    a[0] += 1
    a[1] += rec.get("gpu__time_duration.sum", 0.0)
    a[2] += rec.get("dram__bytes_read.sum", 0.0)
    a[3] += rec.get("dram__bytes_write.sum", 0.0)
print(f"ncu launch list `{sys.argv[1]}` (--clock-control none; per-launch times are serialised and cold-cache)\n")
print("| kernel | launches | mean time | mean dram read | mean dram write |")
print("|---|---|---|---|---|")
for name, (n, t, r, w) in agg.items():
    print(f"| `{name}` | {n} | {t / n / 1e6:.3f} ms | {r / n / 1e9:.3f} GB | {w / n / 1e9:.3f} GB |")
