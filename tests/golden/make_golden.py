#!/usr/bin/env python
"""Generates the committed golden vectors from the REFERENCE ITSELF, run in the build container:

* newton_*.npz   -- per-step vectors out of oracle/_ref/libdaeref.so (the unmodified dae-cpp +
                    Eigen headers behind oracle/ref_shim.cpp): compressed-column Jb of
                    `J.convert(n); Jb -= M_*alpha` (sparse-matrix.hpp:124-152, solver.hpp:523),
                    the BDF derivative (solver.hpp:141-217), b = M*dxdt - f (solver.hpp:516-518)
                    and dx = SparseLU(Jb).solve(b) (solver.hpp:537,558).
* traj_*.json/.f64 -- final state, accepted steps and functor call counts of whole daecpp::solve()
                    runs of hostcpp/_bin/traj_dump_ref (the same driver source compiled against
                    the unmodified reference headers).

Run from the repo root after `python __graft_entry__.py` (needs /root/reference):
    python tests/golden/make_golden.py
The GPU box never has /root/reference; tests read only the files written here.
"""
import json
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as O  # noqa: E402
import problems as P  # noqa: E402

TRAJ = {
    # name: (argv, rtol/atol of the run)
    "flame": (["flame"], 1e-6),
    "perovskite": (["perovskite"], 1e-6),
    "perovskite_shape500": (["perovskite", "shape=1", "n=500"], 1e-6),
    "flame1d_20000": (["flame1d", "n=20000", "t_end=5"], 1e-6),
    "flame1d_shape2000": (["flame1d", "n=2000", "t_end=5", "shape=1"], 1e-6),
    "grid2d_64": (["grid2d", "n=64", "t_end=2"], 1e-6),
}


def newton_case(name, n, J, M, x_hist, xk, f, dt, order):
    assert O.ref is not None, "oracle/_ref/libdaeref.so missing: run `make -C oracle` where /root/reference exists"
    w = np.zeros(5)
    alpha = O.port.oracle_bdf_coefficients(order, np.ascontiguousarray(dt), w)
    dxdt = np.zeros(n)
    hist = np.ascontiguousarray(np.stack(x_hist))
    alpha_ref = O.ref.daeref_time_derivative(n, order, np.ascontiguousarray(dt), np.ascontiguousarray(xk), hist, dxdt)
    assert alpha == alpha_ref
    outer, inner, vals = O.assemble_jb(O.ref, n, J, M, alpha_ref)
    b = np.zeros(n)
    MA, Mi, Mj = M
    O.ref.daeref_residual(n, len(MA), np.ascontiguousarray(MA, np.float64), np.ascontiguousarray(Mi, np.int32),
                          np.ascontiguousarray(Mj, np.int32), dxdt, np.ascontiguousarray(f), b)
    rc, dx, _, _ = O.ref_sparselu(n, (outer, inner, vals), b)
    assert rc == 0
    np.savez_compressed(os.path.join(HERE, f"newton_{name}.npz"), n=n, JA=J[0], Ji=J[1], Jj=J[2], MA=M[0], Mi=M[1], Mj=M[2],
                        hist=hist, xk=xk, f=f, dt=dt, order=order, alpha=alpha_ref, dxdt=dxdt, outer=outer, inner=inner,
                        vals=vals, b=b, dx=dx)
    print(f"newton_{name}: n={n} nnz={len(inner)} alpha={alpha_ref}")


def main():
    r = np.random.default_rng(2024)
    # C2-shaped, small: perovskite N=300, BDF-3 history
    N = 300
    n = 2 * N
    xs = [P.perovskite_state(N, seed=s) for s in range(5)]
    newton_case("perovskite300", n, P.perovskite_jacobian(xs[0], N), P.perovskite_mass(N), xs[1:], xs[0],
                np.linspace(-1, 1, n), np.array([0.01, 0.012, 0.02, 0.0]), 3)
    # C3-shaped, small: tridiagonal n=5000, BDF-1
    n = 5000
    x0 = P.flame1d_state(n)
    xk = x0 + 1e-3 * r.standard_normal(n)
    newton_case("flame1d5000", n, P.flame1d_jacobian(xk), P.identity_mass(n), [x0, x0 * 0, x0 * 0, x0 * 0], xk,
                P.flame1d_rhs(xk), np.array([0.05, 0.0, 0.0, 0.0]), 1)
    # C4-shaped, small: 24 x 24 grid, BDF-2
    m = 24
    n = m * m
    x0 = P.grid2d_state(m)
    xk = x0 + 1e-3 * r.standard_normal(n)
    x1 = x0 - 1e-3 * r.standard_normal(n)
    newton_case("grid2d24", n, P.grid2d_jacobian(xk, m), P.identity_mass(n), [x0, x1, x0 * 0, x0 * 0], xk,
                r.standard_normal(n), np.array([0.02, 0.03, 0.0, 0.0]), 2)
    # random unsorted triplets with duplicates, mass entries outside the Jacobian pattern
    n = 400
    nT = 3000
    Ji = r.integers(0, n, nT).astype(np.int32)
    Jj = r.integers(0, n, nT).astype(np.int32)
    JA = r.standard_normal(nT)
    d = np.arange(n, dtype=np.int32)
    J = (np.concatenate([JA, 50.0 + r.random(n)]), np.concatenate([Ji, d]), np.concatenate([Jj, d]))
    Mi = r.integers(0, n, 500).astype(np.int32)
    Mj = r.integers(0, n, 500).astype(np.int32)
    M = (r.standard_normal(500) * 0.1, Mi, Mj)
    xs = [r.standard_normal(n) for _ in range(5)]
    newton_case("random400", n, J, M, xs[1:], xs[0], r.standard_normal(n), np.array([0.1, 0.2, 0.15, 0.3]), 4)

    exe = os.path.join(ROOT, "hostcpp", "_bin", "traj_dump_ref")
    for name, (argv, tol) in TRAJ.items():
        out = os.path.join(HERE, f"traj_{name}.f64")
        p = subprocess.run([exe] + argv + [f"out={out}"], stdout=subprocess.PIPE, text=True)
        meta = json.loads(p.stdout.strip().splitlines()[-1])
        meta.update(argv=argv, tol=tol)
        for k in ("wall_ms", "ms_rhs_functor", "ms_jac_functor"):
            meta.pop(k, None)
        json.dump(meta, open(os.path.join(HERE, f"traj_{name}.json"), "w"), indent=1)
        print(f"traj_{name}: steps={meta['steps']} status={meta['status']}")


if __name__ == "__main__":
    main()
