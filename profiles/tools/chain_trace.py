"""Debug: per-CTA %globaltimer stamps of the chain kernels on C3 (option trace_units), summarised per kernel."""
import sys, numpy as np
import os
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import importlib
B = importlib.import_module("dae-cpp_b200.binding")
import bench
w = bench.make_workload("flame1d", 1_000_000)
lib = B.load(); c = B.Context(w["n"], 0, lib)
c.set_state(w["x0"]); c.set_mass(w["M"]); c.set_jacobian(w["J"])
alpha = c.residual(w["f"], w["order"], w["dt"]); c.refactor(alpha); c.solve_update(1e-6, 1e-6)
# usage: chain_trace.py [key=value ...]   (daecpp_b200_set_option before the first iteration)
# usage: chain_trace.py [block=K] [key=value ...]   block=K: the traced iterations run as one daecpp_b200_iterations_resident(K) (pipelined)
block = 0
for kv in sys.argv[1:]:
    k, v = kv.split("=", 1)
    if k == "block": block = int(v)
    else: c.set_option(k, float(v))
for _ in range(5): c.iteration_resident(1, w["order"], w["dt"], 1e-6, 1e-6)
c.set_option("trace_units", 1)
if block: c.iterations_resident(block, 1, w["order"], w["dt"], 1e-6, 1e-6)
else:
    for _ in range(3): c.iteration_resident(1, w["order"], w["dt"], 1e-6, 1e-6)
tr = c.unit_trace().astype(np.int64).ravel()
nt = int(c.stat("chain_tiles"))
tr = tr.reshape(5, nt, 8)
names = ["factor tiles", "factor top", "forward tiles", "top solve", "backward tiles"]
t0 = tr[tr > 0].min()
for k in range(5):
    a = tr[k]
    live = a[:, 0] > 0
    if not live.any(): print(names[k], "no stamps"); continue
    a = a[live]
    rel = np.where(a > 0, a - t0, -1) / 1e3
    print(f"== {names[k]} ({live.sum()} CTAs) us since first stamp; slots: 5=entry 0=after wait 1=staged 2=phase0 3=phase1 4=done")
    for slot in (5, 0, 1, 2, 3, 4):
        col = rel[:, slot]; col = col[col >= 0]
        if len(col): print(f"  slot {slot}: min {col.min():7.2f} med {np.median(col):7.2f} max {col.max():7.2f}")
    if a.shape[0] > 3:
        d = np.diff(a[:, [0, 1, 2, 3, 4]], axis=1) / 1e3
        print("  per-CTA phase durations (median): stage %.2f phase0 %.2f phase1 %.2f phase2+out %.2f" % tuple(np.median(d, axis=0)))

a = tr[0]; rel = (a - t0) / 1e3
order_ = np.argsort(rel[:, 4])
print("factor tiles, latest 12 by done time: (tile, staged, done)", [(int(i), round(float(rel[i, 1]), 1), round(float(rel[i, 4]), 1)) for i in order_[-12:]])
print("factor tiles done percentiles 50/90/95/99/100:", np.percentile(rel[:, 4], [50, 90, 95, 99, 100]).round(1))
print("factor tiles staged percentiles 10/50/90/95/99/100:", np.percentile(rel[:, 1], [10, 50, 90, 95, 99, 100]).round(1))
for k, nm in ((2, "forward"), (4, "backward")):
    a = tr[k]; rel = (a - t0) / 1e3; last = 4 if k == 2 else 3
    o = np.argsort(rel[:, last])
    print(nm, "latest 6 (tile, done):", [(int(i), round(float(rel[i, last]), 1)) for i in o[-6:]], "percentiles 50/90/99/100", np.percentile(rel[:, last], [50, 90, 99, 100]).round(1))

for k, nm in ((3, "top solve"), (4, "backward tiles")):
    a = tr[k]; live = a[:, 0] > 0; a = a[live]
    end = 4 if k == 3 else 3; st = 3 if k == 3 else 2
    print(nm, "backward part (median us): levels 2+1 %.2f, level 0 sweep %.2f, write-out %.2f" % (
        np.median(a[:, 6] - a[:, st]) / 1e3, np.median(a[:, 7] - a[:, 6]) / 1e3, np.median(a[:, end] - a[:, 7]) / 1e3))

a0 = tr[0, 0]
if a0[6] > 0 and a0[7] > 0:
    print("tile 0: previous iteration factor entry -> backward done %.2f us; idle until this iteration's factor entry %.2f us; period %.2f us" % (
        (a0[7] - a0[6]) / 1e3, (a0[5] - a0[7]) / 1e3, (a0[5] - a0[6]) / 1e3))
