"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel name.
usage: python tools/launch_summary.py file.csv [first last]  (optional window to print in order)"""
import collections
import csv
import re
import sys

with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith("==")]
rows = list(csv.DictReader(lines))


def us(row):
    v = float(row["Metric Value"].replace(",", ""))
    return {"ns": v / 1e3, "us": v, "ms": v * 1e3}.get(row["Metric Unit"], v)


def short(n):
    return re.sub(r"\(.*", "", n).replace("void ", "").replace("dtg::", "")[:48]


agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    a = agg[short(r["Kernel Name"])]
    a[0] += 1
    a[1] += us(r)
tot = sum(v[1] for v in agg.values())
print("total %.0f us over %d launches" % (tot, len(rows)))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print("%9.0f us %5.1f%% n=%4d avg %7.1f  %s" % (v[1], 100 * v[1] / tot, v[0], v[1] / v[0], k))
if len(sys.argv) > 3:
    for i, r in enumerate(rows[int(sys.argv[2]):int(sys.argv[3])]):
        print(i + int(sys.argv[2]), short(r["Kernel Name"]), "%.1f" % us(r), r["Grid Size"])
