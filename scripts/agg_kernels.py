"""Aggregate an `ncu --metrics gpu__time_duration.sum --csv` log by kernel name."""
import collections
import csv
import re
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
with open(path) as f:
    lines = [l for l in f if l.startswith('"')]
agg = collections.defaultdict(lambda: [0, 0.0])
tot = 0
rows = []
for row in csv.DictReader(lines):
    v = float(row['Metric Value'].replace(',', ''))
    unit = row['Metric Unit']
    v = v / 1e3 if unit == 'ns' else v * 1e3 if unit == 'ms' else v
    short = re.sub(r'\(.*', '', row['Kernel Name'])[:100]
    agg[short][0] += 1
    agg[short][1] += v
    tot += v
    rows.append((short, v, row.get('Grid Size'), row.get('Block Size')))
print(f"total {tot / 1e3:.3f} ms over {len(rows)} kernels")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{t:9.1f} us {c:4d} avg {t / c:6.1f}  {k}")
print("--- slowest individual")
for s, v, g, b in sorted(rows, key=lambda x: -x[1])[:15]:
    print(f"{v:8.1f} {g} {b} {s[:80]}")
