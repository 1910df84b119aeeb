"""Per-kernel means of an `ncu --metrics a,b,c --csv` log (one row per kernel, one column per metric, plus the share of time).
Usage: python tools/ncu_kernels.py log.csv"""
import collections, csv, sys
lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
rows = list(csv.DictReader(lines))
val = collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows:
    k = r["Kernel Name"].split("(")[0].replace("void adroit::", "").replace("void ", "")[:48]
    try:
        v = float(r["Metric Value"].replace(",", ""))
    except ValueError:
        continue
    val[k][r["Metric Name"]].append(v)
metrics = sorted({m for k in val for m in val[k]})
short = [m.replace("smsp__", "").replace("sm__", "").replace("launch__", "").replace(".avg.pct_of_peak_sustained_active", "%").replace(".sum", "")[:22] for m in metrics]
tot = sum(sum(v.get("gpu__time_duration.sum", [0])) for v in val.values())
print(f"{'kernel':48s} {'n':>5s} {'share':>6s} " + " ".join(f"{s:>22s}" for s in short))
for k, v in sorted(val.items(), key=lambda kv: -sum(kv[1].get("gpu__time_duration.sum", [0]))):
    t = sum(v.get("gpu__time_duration.sum", [0]))
    n = len(next(iter(v.values())))
    print(f"{k:48s} {n:5d} {100*t/max(tot,1e-9):5.1f}% " + " ".join(f"{sum(v[m])/max(len(v[m]),1):22.1f}" if m in v else f"{'':>22s}" for m in metrics))
