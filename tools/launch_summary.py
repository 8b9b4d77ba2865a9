"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name.
    python tools/launch_summary.py gpurun_out/launches_eval.csv [out.txt]"""
import csv
import sys
from collections import defaultdict

rows = []
with open(sys.argv[1]) as f:
    lines = [ln for ln in f if ln.startswith('"')]
rd = csv.DictReader(lines)
agg = defaultdict(lambda: [0, 0.0])
order = []
for r in rd:
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    name = r["Kernel Name"].split("(")[0].replace("<unnamed>::", "").replace("void ", "")
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    us = v / 1e3 if unit.startswith("ns") else (v * 1e3 if unit.startswith("ms") else v)
    agg[name][0] += 1
    agg[name][1] += us
    order.append((name, us))
tot = sum(v[1] for v in agg.values())
out = [f"{'kernel':60s} {'calls':>6s} {'total_us':>10s} {'avg_us':>8s} {'share':>6s}"]
for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    out.append(f"{name[:60]:60s} {n:6d} {us:10.1f} {us / n:8.2f} {100 * us / tot:5.1f}%")
out.append(f"{'TOTAL':60s} {sum(v[0] for v in agg.values()):6d} {tot:10.1f}")
txt = "\n".join(out) + "\n"
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(txt)
print(txt)
