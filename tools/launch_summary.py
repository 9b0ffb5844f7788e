"""Per-kernel totals of an ncu launch list (gpu__time_duration.sum CSV): python tools/launch_summary.py file.csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
h = rows[hdr]
ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    k = r[ki].split("(")[0][:60]
    v = float(r[vi].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
    agg.setdefault(k, [0, 0.0])
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v for _, v in agg.values())
print(f"total {tot:.3f} ms over {sum(n for n, _ in agg.values())} launches")
for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:20]:
    print(f"{v:9.3f} ms {n:4d} {100 * v / tot:5.1f}% {k}")
