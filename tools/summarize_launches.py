"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launches, total and share."""
import csv, sys, collections, re
path = sys.argv[1]
rows = []
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
rd = csv.DictReader(lines)
agg = collections.OrderedDict()
tot = 0.0
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    us = v / 1e3 if u in ("ns", "nsecond") else v * (1e3 if u in ("ms", "msecond") else 1.0)
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    name = re.sub(r"^void ", "", name)
    grid = r.get("Grid Size", "")
    key = name
    a = agg.setdefault(key, [0, 0.0])
    a[0] += 1
    a[1] += us
    tot += us
print(f"{'kernel':70s} {'launches':>8s} {'total_us':>12s} {'avg_us':>9s} {'share':>7s}")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:70]:70s} {n:8d} {t:12.1f} {t/n:9.1f} {100*t/tot:6.1f}%")
print(f"{'TOTAL':70s} {sum(a[0] for a in agg.values()):8d} {tot:12.1f}")
