"""The Newton linear step at the sizes BASELINE.json names (C3: flame-1D n = 10^6; C4: 2-D grid 1024^2), through
the default device path, against figures the REFERENCE produced for the very same inputs
(tests/golden/make_golden_large.py -> tests/golden/solve_*.json; Eigen::SparseLU needs 40 s for the grid, so
its residual norm and a strided sample of its dx are committed instead of being recomputed on the GPU box).
Bar: assembly bit-exact; backward error <= 1e-12; ||b - Jb dx||_2 <= 10 ||b - Jb dx_ref||_2 + 1e-12 ||b||_2
(SURVEY.md 8(c) definitions 1 and 2; BASELINE.json north_star)."""
import json
import os

import numpy as np
import pytest

import oracle_lib as O
from gpu_lib import binding as B

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["c3_flame1d_1e6", "c4_grid2d_1024"])
def test_newton_solve_at_baseline_size_matches_reference(name):
    import bench
    meta = json.load(open(os.path.join(GOLD, f"solve_{name}.json")))
    w = bench.make_workload(meta["workload"], meta["size"])
    n, dt = w["n"], w["dt"]
    assert n == meta["n"]
    c = B.Context(n)
    c.set_state(w["x0"]); c.set_mass(w["M"]); c.set_jacobian(w["J"])
    alpha = c.residual(w["f"], 1, dt)
    assert alpha == meta["alpha"]
    rc, flags = c.iteration_resident(1, 1, dt, 1e-6, 1e-6)
    assert rc == 0
    assert c.stat("n_symbolic") == 1
    dx = c.get_vector("dx")
    got_csc = c.get_csc(2)
    assert c.get_vector("xk").tobytes() == (w["x0"] + dx).tobytes()
    c.close()
    csc = O.assemble_jb(O.port, n, w["J"], w["M"], alpha)
    assert len(csc[1]) == meta["nnz_jb"]
    for got, want in zip(got_csc, csc):
        assert got.tobytes() == want.tobytes(), "Jb = J - alpha*M is not bit-exact"
    b = -np.ascontiguousarray(w["f"])  # xk == x[0]: dxdt = 0, b = M*0 - f
    assert abs(float(np.linalg.norm(b)) - meta["b_norm2"]) <= 1e-12 * meta["b_norm2"]  # same inputs as the reference saw
    eta, r2 = O.backward_error(n, csc, dx, b)
    assert eta <= 1e-12, eta
    assert r2 <= 10.0 * meta["r2_ref"] + 1e-12 * meta["b_norm2"], (r2, meta["r2_ref"])
    ref = np.array(meta["dx_ref_sample"])
    assert np.max(np.abs(dx[::meta["stride"]] - ref)) <= 1e-8 * meta["dx_ref_maxabs"]
