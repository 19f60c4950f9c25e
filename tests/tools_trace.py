"""Developer tool (not a test): per-unit timeline of the factor kernel.  python tests/tools_trace.py [workload] [n]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gpu_lib import binding as B
import bench

name = sys.argv[1] if len(sys.argv) > 1 else "flame1d"
size = int(sys.argv[2]) if len(sys.argv) > 2 else 0
w = bench.make_workload(name, size)
c = B.Context(w["n"])
for kv in [a for a in sys.argv[3:] if "=" in a]:
    k, v = kv.split("=")
    c.set_option(k, float(v))
c.set_state(w["x0"]); c.set_mass(w["M"]); c.set_jacobian(w["J"])
a = c.residual(w["f"], 1, w["dt"])
c.refactor(a)
c.set_option("trace_units", 1)
for _ in range(3):
    c.refactor(a)
print("refactor ms", c.stat("ms_last_refactor"), "solve ms (next)")
c.solve_update(1e-6, 1e-6)
print("solve ms", c.stat("ms_last_solve"))
nu = int(c.stat("n_units"))
T = c.unit_trace().astype(np.float64)[:3 * nu].reshape(3, -1, 4)  # (the rows behind them: per-level stamps of one unit)
print({k: c.stat(k) for k in ("n_units", "n_stages", "factor_smem_bytes", "solve_smem_bytes", "unit_grid_factor", "unit_grid_solve", "n_fronts")})
for name, t in zip(("factor", "forward", "backward"), T):
    t = (t - t[:, 0].min()) / 1e3
    d = t[:, 3] - t[:, 0]
    print("%-8s span %.1f us | unit: mean %.2f p50 %.2f max %.2f | stage %.2f wait %.2f levels %.2f" % (
        name, t[:, 3].max(), d.mean(), np.median(d), d.max(), (t[:, 1] - t[:, 0]).mean(), (t[:, 2] - t[:, 1]).mean(),
        (t[:, 3] - t[:, 2]).mean()))
    if "-v" in sys.argv:
        n = len(t)
        for k in list(range(0, n, max(1, n // 8))) + list(range(max(0, n - 12), n)):
            print("   ", k, "start %.1f staged %.1f dep %.1f done %.1f" % tuple(t[k]))
