"""sum gpu__time_duration per kernel name from an ncu --csv launch list; optional: only the last K launches"""
import csv, sys, collections, re
rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") == "gpu__time_duration.sum":
        v = float(r["Metric Value"].replace(",", "")); u = r.get("Metric Unit", "ns")
        v *= {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "nsecond": 1e-3, "ms": 1e3, "msecond": 1e3}.get(u, 1e-3)
        rows.append((re.sub(r"\(.*", "", r["Kernel Name"]), v))
last = int(sys.argv[2]) if len(sys.argv) > 2 else len(rows)
rows = rows[-last:]
tot = collections.OrderedDict()
for k, v in rows:
    a = tot.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
all_us = sum(v for _, v in rows)
for k, (n, v) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
    print(f"{k:40s} {n:5d} launches {v:10.1f} us  {100 * v / all_us:5.1f} %   avg {v / n:8.1f} us")
print(f"total {all_us:.1f} us over {len(rows)} launches")
