"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list by kernel.
Usage: python tools/ncu_launches.py launches.csv"""
import collections, csv, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
rows = list(csv.DictReader(lines))
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    k = r["Kernel Name"].split("(")[0][:80]
    agg[k][0] += 1
    agg[k][1] += float(r["Metric Value"].replace(",", "")) * (1e3 if r["Metric Unit"].startswith("us") else 1.0)
tot = sum(v[1] for v in agg.values())
print(f"{'launches':>8} {'total_us':>12} {'share':>7} {'avg_us':>10}  kernel")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{v[0]:8d} {v[1]/1e3:12.1f} {100*v[1]/tot:6.2f}% {v[1]/1e3/v[0]:10.1f}  {k}")
