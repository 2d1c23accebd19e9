"""Summarise an ncu `--metrics gpu__time_duration.sum --csv` launch list by kernel name.
    python tools/summarize_launches.py gpurun_out/launches.csv > profiles/<round>_launches_summary.txt"""
import collections
import csv
import re
import sys

rows = list(csv.DictReader(l for l in open(sys.argv[1]) if not l.startswith("==")))
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    v = float(r["Metric Value"].replace(",", ""))
    v = v / 1000 if r["Metric Unit"] == "ns" else v * 1000 if r["Metric Unit"] == "ms" else v
    name = re.sub(r"\(.*", "", re.sub(r"<.*", "", r["Kernel Name"]))[:64]
    agg[name][0] += 1
    agg[name][1] += v
tot = sum(v[1] for v in agg.values())
print(f"# {len(rows)} launches, {tot:.1f} us total (ncu-serialised, cold cache: compare SHARES, not absolutes)")
print(f"# {'total_us':>10} {'count':>6} {'avg_us':>8} {'share':>6}  kernel")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"  {v[1]:10.1f} {v[0]:6d} {v[1] / v[0]:8.2f} {100 * v[1] / tot:5.1f}%  {k}")
