"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel.
    python tools/summarize_launches.py gpurun_out/launches_bench.csv [steps_in_capture] > profiles/...md
"""
import collections
import csv
import re
import sys

path = sys.argv[1]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
lines = [l for l in open(path) if not l.startswith("==")]
agg = collections.defaultdict(lambda: [0, 0.0])
n = 0
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    u = r["Metric Unit"]
    v = v / 1e3 if u == "ns" else (v * 1e3 if u == "ms" else v)
    name = re.sub(r"<.*", "", r["Kernel Name"])
    name = re.sub(r"^void ", "", name)[:90]
    agg[name][0] += 1
    agg[name][1] += v
    n += 1
tot = sum(v[1] for v in agg.values())
print(f"launches: {n} ({n / steps:.0f} per step over {steps} steps); total kernel time {tot / 1e3:.2f} ms "
      f"({tot / 1e3 / steps:.2f} ms per step; ncu-serialised, cold cache: compare SHARES)\n")
print("| share | total us | launches | avg us | kernel |")
print("|---:|---:|---:|---:|---|")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    print(f"| {v[1] / tot * 100:.2f}% | {v[1]:.1f} | {v[0]} | {v[1] / v[0]:.1f} | `{k}` |")
