"""Aggregate an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel name for one step.
usage: summarise_launches.py launches.csv <launches per step> <step index>"""
import collections
import csv
import re
import sys

path, per, idx = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
lines = [l for l in open(path) if not l.startswith("==")]
rows = list(csv.DictReader(lines))
seg = rows[per * idx:per * (idx + 1)]
agg = collections.defaultdict(lambda: [0, 0.0])
for r in seg:
    name = re.sub(r"\(.*", "", r["Kernel Name"])
    v = float(r["Metric Value"].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(r["Metric Unit"], 1e-6)
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print(f"| kernel | launches | ms (serialised, cold) | share |\n|---|---|---|---|")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"| `{k[:90]}` | {v[0]} | {v[1]:.3f} | {100 * v[1] / tot:.1f}% |")
print(f"| **total** | {sum(v[0] for v in agg.values())} | {tot:.3f} | |")
