"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel."""
import collections
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    try:
        v = float(r[vi].replace(",", ""))
    except ValueError:
        continue
    k = r[ki][:70]
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v[1] for v in agg.values())
print(f"{'kernel':70s} {'n':>5s} {'total ms':>10s} {'avg us':>10s} {'share':>6s}")
for k, (c, v) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{k:70s} {c:5d} {v / 1e6:10.3f} {v / c / 1e3:10.1f} {v / tot:6.3f}")
