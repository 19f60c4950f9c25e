"""Tiny systems (n <= 32): the one-launch Newton iteration of dae-cpp_b200/csrc/tiny.cu (mapped hand-off, dense LU
with partial pivoting on the device) behind daecpp_b200_iteration_host, against the oracle:
b and Jb bit for bit (duplicates summed in input order, mass entries off the Jacobian pattern, singular mass),
dx by backward error and against Eigen, flags, history / accept / rollback interplay with the regular entry
points, decomposition failure, and the pattern cache.  The reference's own test suite (mostly N <= 4) runs through
this path in tests/test_golden.py::test_reference_gtest_suite_against_overlay."""
import numpy as np
import pytest

import oracle_lib as O
import problems as P
from gpu_lib import binding as B

pytestmark = pytest.mark.gpu


def _random_system(n, seed, dup=True):
    r = np.random.default_rng(seed)
    nT = 3 * n + (5 if dup else 0)
    Ji = r.integers(0, n, nT).astype(np.int32); Jj = r.integers(0, n, nT).astype(np.int32)
    JA = r.standard_normal(nT)
    d = np.arange(n, dtype=np.int32)
    J = (np.concatenate([JA, 4.0 + r.random(n)]), np.concatenate([Ji, d]), np.concatenate([Jj, d]))
    nM = max(1, n // 2) + 2
    M = (r.standard_normal(nM), r.integers(0, n, nM).astype(np.int32), r.integers(0, n, nM).astype(np.int32))
    return J, M


@pytest.mark.parametrize("n,order", [(1, 1), (2, 2), (3, 3), (4, 4), (7, 2), (32, 4)])
def test_tiny_iteration_matches_oracle(n, order):
    J, M = _random_system(n, 100 + n)
    r = np.random.default_rng(n)
    xs = [r.standard_normal(n) for _ in range(5)]
    f = r.standard_normal(n)
    dt = np.array([0.1, 0.2, 0.15, 0.3])
    c = B.Context(n)
    c.set_state(xs[0])
    for k in range(4):
        c.set_vector(f"x{k}", xs[k + 1])
    c.set_mass(M); c.set_jacobian(J)
    xk_out = np.zeros(n)
    rc, flags, alpha = c.iteration_host(f, 1, order, dt, 1e-6, 1e-6, xk_out)
    assert rc == 0
    assert c.stat("kernels_launched") <= 3  # the iteration itself is ONE launch (plus set_state's copies)
    # oracle: residual and Jb bit for bit
    w = np.zeros(5)
    assert alpha == O.port.oracle_bdf_coefficients(order, dt, w)
    dxdt = np.zeros(n)
    O.port.oracle_time_derivative(n, order, dt, np.ascontiguousarray(xs[0]), np.ascontiguousarray(np.stack(xs[1:]).reshape(-1)), dxdt)
    b = np.zeros(n)
    O.port.oracle_residual(n, *O.convert(O.port, n, M), dxdt, f, b)
    assert c.get_vector("b").tobytes() == b.tobytes()
    csc = O.assemble_jb(O.port, n, J, M, alpha)
    assert tuple(a.tobytes() for a in c.get_csc(2)) == tuple(a.tobytes() for a in csc)
    dx = c.get_vector("dx")
    eta, r2 = O.backward_error(n, csc, dx, b)
    assert eta <= 1e-12
    if O.ref is not None:
        rc_ref, dx_ref, _, _ = O.ref_sparselu(n, csc, b)
        assert rc_ref == 0 and np.max(np.abs(dx - dx_ref)) <= 1e-9 * max(1.0, np.max(np.abs(dx_ref)))
    xk = xs[0].copy()
    want_flags = O.port.oracle_error_check_update(n, dx, np.ascontiguousarray(xs[1]), xk, 1e-6, 1e-6, 1e100)
    assert flags == want_flags
    assert xk_out.tobytes() == xk.tobytes() == c.get_vector("xk").tobytes()
    # a second iteration on the same factors, after an accepted step: history shift, speculative variability
    v = c.accept_step(True, 1e-4)
    hist = np.stack(xs[1:]).reshape(-1).copy()
    assert v == O.port.oracle_accept_step(n, xk, hist, 1, 1e-4)
    for k in range(4):
        assert c.get_vector(f"x{k}").tobytes() == hist[k * n:(k + 1) * n].tobytes()
    rc, flags2, _ = c.iteration_host(f * 0.5, 0, order, dt, 1e-6, 1e-6, xk_out)
    assert rc == 0
    O.port.oracle_time_derivative(n, order, dt, xk, hist, dxdt)
    O.port.oracle_residual(n, *O.convert(O.port, n, M), dxdt, f * 0.5, b)
    assert c.get_vector("b").tobytes() == b.tobytes()
    eta, _ = O.backward_error(n, csc, c.get_vector("dx"), b)
    assert eta <= 1e-12
    # rollback restores the accepted state
    c.rollback()
    assert c.get_vector("xk").tobytes() == xk.tobytes()
    c.close()


def test_tiny_and_regular_paths_agree():
    n = 6
    J, M = _random_system(n, 9, dup=False)
    r = np.random.default_rng(1)
    x0, f = r.standard_normal(n), r.standard_normal(n)
    dt = np.array([0.05, 0, 0, 0.0])
    out = []
    for tiny in (1, 0):
        c = B.Context(n)
        c.set_option("tiny", tiny)
        c.set_state(x0); c.set_mass(M); c.set_jacobian(J)
        xk = np.zeros(n)
        rc, flags, alpha = c.iteration_host(f, 1, 1, dt, 1e-6, 1e-6, xk)
        out.append((rc, flags, alpha, c.get_vector("b").tobytes(), tuple(a.tobytes() for a in c.get_csc(2)), c.get_vector("dx")))
        c.close()
    assert out[0][:5] == out[1][:5]
    assert np.max(np.abs(out[0][5] - out[1][5])) <= 1e-12 * np.max(np.abs(out[1][5]))


def test_tiny_singular_matrix_and_bad_index_and_pattern_cache():
    c = B.Context(2)
    c.set_state(np.zeros(2))
    c.set_mass((np.zeros(0), np.zeros(0, np.int32), np.zeros(0, np.int32)))
    J = (np.array([1.0, 2.0, 2.0, 4.0]), np.array([0, 0, 1, 1], np.int32), np.array([0, 1, 0, 1], np.int32))
    c.set_jacobian(J)
    xk = np.zeros(2)
    rc, _, _ = c.iteration_host(np.ones(2), 1, 1, np.array([0.1, 0, 0, 0.0]), 1e-6, 1e-6, xk, allow=(B.EDECOMP,))
    assert rc == B.EDECOMP  # rank one: Eigen reports the same (linsolver_failed_decomposition, solver.hpp:541-551)
    rc, _, _ = c.iteration_host(np.ones(2), 0, 1, np.array([0.1, 0, 0, 0.0]), 1e-6, 1e-6, xk, allow=(B.EINVAL,))
    assert rc == B.EINVAL  # no valid factorisation to reuse
    assert c.set_jacobian((np.ones(1), np.array([2], np.int32), np.array([0], np.int32)), allow=(B.EINDEX,)) == B.EINDEX
    # values-only hand-off on the cached stream
    c.set_jacobian((np.array([3.0, 0.0, 0.0, 5.0]), J[1], J[2]))
    lib = c.lib
    vals = np.array([2.0, 0.0, 0.0, 8.0])
    assert lib.daecpp_b200_set_jacobian_values(c.h, vals.ctypes.data, 4) == 0
    c.set_state(np.zeros(2))  # (the failed iteration left garbage in the iterate: callers roll back, solver.hpp:541-551)
    rc, flags, _ = c.iteration_host(np.array([2.0, 8.0]), 1, 1, np.array([0.1, 0, 0, 0.0]), 1e-6, 1e-6, xk)
    assert rc == 0 and np.allclose(xk, [-1.0, -1.0])  # J dx = -f
    c.close()
