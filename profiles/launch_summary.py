"""Aggregates an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel (profiles/ helper).
usage: python profiles/launch_summary.py gpurun_out/launches.csv > profiles/rNN_launches.md"""
import collections
import csv
import re
import sys

lines = open(sys.argv[1]).read().splitlines()
hi = next(i for i, l in enumerate(lines) if l.startswith('"ID"'))
rows = list(csv.DictReader(lines[hi:]))
agg = collections.OrderedDict()
for r in rows:
    n = re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "")
    a = agg.setdefault(n, [0, 0.0, r["Grid Size"], r["Block Size"]])
    a[0] += 1
    a[1] += float(r["Metric Value"])
tot = sum(v[1] for v in agg.values())
print(f"launches: {len(rows)}; total kernel time {tot / 1e3:.1f} us (ncu-serialised, cold caches: shares matter, not absolutes)\n")
print("| kernel | launches | total us | us/launch | share | grid | block |")
print("|---|---|---|---|---|---|---|")
for n, (c, t, g, b) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{n[:90]}` | {c} | {t / 1e3:.1f} | {t / 1e3 / c:.1f} | {t / tot * 100:.1f}% | {g} | {b} |")
