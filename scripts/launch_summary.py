#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv):  python scripts/launch_summary.py FILE [top N]"""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 14
h = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[h]; ik = hdr.index("Kernel Name"); iv = hdr.index("Metric Value"); iu = hdr.index("Metric Unit")
t = collections.Counter(); c = collections.Counter()
for r in rows[h + 1:]:
    if len(r) <= iv:
        continue
    k = r[ik].split("(")[0][-44:]
    v = float(r[iv].replace(",", "")) * {"us": 1e3, "usecond": 1e3, "ms": 1e6, "msecond": 1e6}.get(r[iu], 1)
    t[k] += v; c[k] += 1
tot = sum(t.values())
for k, v in t.most_common(top):
    print(f"{k:46s} {c[k]:5d} launches {v / 1e6:8.2f} ms {100 * v / tot:5.1f}%")
print("total", round(tot / 1e6, 2), "ms over", sum(c.values()), "launches")
