#!/usr/bin/env python
"""Per-kernel and per-level summary of an ncu launch list of the C4 bench command
(ncu --metrics gpu__time_duration.sum --clock-control none --csv ... python bench.py --workload grid2d ...).
Usage: python tools/c4_levels.py <launches.csv> [iterations to keep from the end, default 1]"""
import csv
import re
import sys
from collections import OrderedDict

rows = []
with open(sys.argv[1]) as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r["Metric Name"] != "gpu__time_duration.sum":
        continue
    name = re.sub(r"\(.*", "", r["Kernel Name"]).split("::")[-1]
    rows.append((name, float(r["Metric Value"]) / 1e3, r["Grid Size"]))
# one Newton iteration = from a k_scatter_big (refactor start) to the launch before the next one
starts = [i for i, r in enumerate(rows) if r[0].startswith("k_scatter_big")]
if len(starts) >= 2:
    rows = rows[starts[-2]:starts[-1]]
    ends = [i for i, r in enumerate(rows) if r[0].startswith("k_permute_out")]
    if ends:  # the bench repeats the solve phase on its own: keep the first one (and the update behind it)
        e = ends[0] + 1
        if e < len(rows) and rows[e][0].startswith("k_update"):
            e += 1
        rows = rows[:e]
per = OrderedDict()
for name, us, grid in rows:
    n, t = per.get(name, (0, 0.0))
    per[name] = (n + 1, t + us)
tot = sum(t for _, t in per.values())
print(f"# one Newton iteration (refactor + solve), {len(rows)} launches, serialised sum {tot / 1e3:.3f} ms")
try:
    sys.stdout.reconfigure(line_buffering=True)
except Exception:
    pass
for name, (n, t) in sorted(per.items(), key=lambda kv: -kv[1][1]):
    print(f"{name:28s} {n:4d} launches {t / 1e3:8.3f} ms {100 * t / tot:5.1f}%  avg {t / n:8.1f} us")
print("\n## k_unit_factor launches (us, grid)")
for name, us, grid in rows:
    if name == "k_unit_factor":
        print(f"  {us:8.1f}  {grid}")
