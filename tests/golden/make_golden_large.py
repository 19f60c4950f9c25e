#!/usr/bin/env python
"""Golden figures for the BASELINE-SIZE configurations, from the REFERENCE ITSELF (oracle/_ref/libdaeref.so =
unmodified dae-cpp + Eigen::SparseLU; hostcpp/_bin/traj_dump_ref), run in the build container where
/root/reference exists.  The vectors are large (8 MB each), so only what the parity bar needs is kept:

* solve_<name>.json -- for the Newton linear system of bench.py's workload (C3: flame-1D n = 10^6, C4: 2-D grid
                       1024^2; first iteration of a BDF-1 step, b = -f): the reference's residual norm
                       ||b - Jb dx_ref||_2, its normwise backward error, ||b||_2, ||dx_ref||_inf, and dx_ref
                       sampled at a fixed stride.  Bar (SURVEY.md 8(c) definition 2):
                       ||b - Jb dx||_2 <= 10 ||b - Jb dx_ref||_2 + 1e-12 ||b||_2.
* traj_flame1d_shape200000.* -- a whole daecpp::solve with 4.8 MB of Jacobian triplet values: the run that makes
                       the overlay page-lock the user's triplet storage (solver.hpp overlay: DeviceNewton::pin).

    python tests/golden/make_golden_large.py        # ~3 minutes (the reference factors the 1024^2 grid in ~65 s)
"""
import json
import os
import subprocess
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)

import oracle_lib as O  # noqa: E402
import bench  # noqa: E402  (make_workload: the very inputs bench.py times)

STRIDE = 997


def solve_case(name, workload, size):
    w = bench.make_workload(workload, size)
    n = w["n"]
    alpha = 1.0 / w["dt"][0]
    csc = O.assemble_jb(O.ref, n, w["J"], w["M"], alpha)
    b = -np.ascontiguousarray(w["f"])
    t0 = time.time()
    rc, dx, _, _ = O.ref_sparselu(n, csc, b)
    assert rc == 0
    eta, r2 = O.backward_error(n, csc, dx, b)
    meta = {"workload": workload, "size": size, "n": n, "alpha": alpha, "nnz_jb": int(len(csc[1])), "r2_ref": r2, "eta_ref": eta,
            "b_norm2": float(np.linalg.norm(b)), "dx_ref_maxabs": float(np.max(np.abs(dx))), "stride": STRIDE,
            "dx_ref_sample": dx[::STRIDE].tolist(), "reference_seconds": time.time() - t0}
    json.dump(meta, open(os.path.join(HERE, f"solve_{name}.json"), "w"))
    print(f"solve_{name}: n={n} r2_ref={r2:.3e} eta_ref={eta:.3e} ({meta['reference_seconds']:.1f} s)")


def traj_case(name, argv, tol):
    exe = os.path.join(ROOT, "hostcpp", "_bin", "traj_dump_ref")
    out = os.path.join(HERE, f"traj_{name}.f64")
    p = subprocess.run([exe] + argv + [f"out={out}"], stdout=subprocess.PIPE, text=True)
    meta = json.loads(p.stdout.strip().splitlines()[-1])
    full = np.fromfile(out)
    os.remove(out)
    full[::STRIDE].tofile(out)  # a strided sample of the final state is enough (and 1.6 kB instead of 1.6 MB)
    meta.update(argv=argv, tol=tol, stride=STRIDE)
    for k in ("wall_ms", "first_wall_ms", "ms_rhs_functor", "ms_jac_functor"):
        meta.pop(k, None)
    json.dump(meta, open(os.path.join(HERE, f"traj_{name}.json"), "w"), indent=1)
    print(f"traj_{name}: steps={meta['steps']} status={meta['status']}")


if __name__ == "__main__":
    traj_case("flame1d_shape200000", ["flame1d", "n=200000", "t_end=5", "shape=1"], 1e-6)
    solve_case("c3_flame1d_1e6", "flame1d", 1_000_000)
    solve_case("c4_grid2d_1024", "grid2d", 1024)
