"""Aggregate an ncu launch list (--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv) by
kernel name: count, summed time, DRAM bytes.  usage: python tools/agg_launches.py launches.csv [skip_launches]"""
import csv, re, sys
from collections import OrderedDict

rows = []
with open(sys.argv[1], newline="") as f:
    lines = [l for l in f if l.startswith('"')]
rd = csv.DictReader(lines)
per = OrderedDict()
for r in rd:
    k = (r["ID"], r["Kernel Name"])
    per.setdefault(k, {})[r["Metric Name"]] = (float(r["Metric Value"].replace(",", "")), r["Metric Unit"])
skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
agg = {}
unit_scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
for (i, name), m in per.items():
    if int(i) < skip:
        continue
    name = re.sub(r"\(.*", "", name).replace("yc::", "").strip()
    a = agg.setdefault(name, [0, 0.0, 0.0, 0.0])
    a[0] += 1
    for key, slot in (("gpu__time_duration.sum", 1), ("dram__bytes_read.sum", 2), ("dram__bytes_write.sum", 3)):
        if key in m:
            v, u = m[key]
            a[slot] += v * unit_scale.get(u, 1.0)
tot = 0.0
for name, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    tot += a[1]
    bw = (a[2] + a[3]) / a[1] * 1e3 if a[1] else 0.0
    print(f"{name[:52]:52s} n={a[0]:4d} t={a[1]:9.1f}us rd={a[2]:9.1f}MB wr={a[3]:9.1f}MB {bw:8.1f} GB/s")
print(f"total {tot:.1f} us over {sum(a[0] for a in agg.values())} launches (ncu: cold caches, serialised)")
