#!/usr/bin/env python
"""Developer tool: where the leaf-subtree factor launch of C4 spends its time.
Per-unit phase stamps (start / entries scattered / children ready / levels done) of every unit of the first launch
group, and the local levels of one unit one by one.  Usage: python tools/unit_trace.py [side] [unit]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import problems as P
from gpu_lib import binding as B

m = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
unit = int(sys.argv[2]) if len(sys.argv) > 2 else 0
n = m * m
u = P.grid2d_state(m)
c = B.Context(n)
c.set_option("chain", 0); c.set_option("trace_units", 1); c.set_option("trace_unit_index", unit)
for kv in sys.argv[3:]:
    k, v = kv.split("="); c.set_option(k, float(v))
c.set_mass(P.identity_mass(n)); c.set_jacobian(P.grid2d_jacobian(u, m))
c.refactor(50.0); c.refactor(50.0)
nu = int(c.stat("n_units"))
tr = c.unit_trace()
fac = tr[:nu]
lev = tr[3 * nu + 1:3 * nu + 1 + 60]
t0 = fac[:, 0][fac[:, 0] > 0].min()
first = fac[:444] if nu >= 444 else fac
d = (first - t0) / 1e3
print(f"units {nu}; first launch group ({len(first)} units), us relative to the first start")
print("  start       min/med/max", np.min(d[:, 0]), np.median(d[:, 0]), np.max(d[:, 0]))
print("  scattered   min/med/max (duration)", *[f(d[:, 1] - d[:, 0]) for f in (np.min, np.median, np.max)])
print("  levels      min/med/max (duration)", *[f(d[:, 3] - d[:, 2]) for f in (np.min, np.median, np.max)])
print("  end         min/med/max", np.min(d[:, 3]), np.median(d[:, 3]), np.max(d[:, 3]))
print(f"unit {unit}: local levels (us since levels began, width, fronts, largest order)")
start = fac[unit, 2]
prev = start
for l in range(60):
    if lev[l, 0] == 0:
        break
    print(f"  level {l:2d}  +{(lev[l, 0] - prev) / 1e3:8.1f} us  G={lev[l, 1]:3d} fronts={lev[l, 2]:4d} max_d={lev[l, 3]:4d}")
    prev = lev[l, 0]
c.close()
