"""Pins the CPU restatement (oracle/oracle.c) against (1) literals from the reference's own
tests and (2) the unmodified reference headers (oracle/_ref/libdaeref.so).  CPU only."""
import numpy as np
import pytest

import oracle_lib as O
import problems as P

needs_ref = pytest.mark.skipif(O.ref is None, reason="oracle/_ref not built")


def test_port_loaded():
    assert O.port is not None, "run `make -C oracle` (or __graft_entry__.build())"


# --- literals from /root/reference/tests/test_sparse-matrix.cpp:183-229 (convert, duplicates)
def test_convert_kat_duplicates_summed():
    # M2 of the reference test: (0,0,-5.0) (1,1,2.5) (1,1,2.5) -> dense [[-5,0],[0,5]]
    A = np.array([-5.0, 2.5, 2.5]); i = np.array([0, 1, 1], np.int32); j = np.array([0, 1, 1], np.int32)
    outer, inner, vals = O.convert(O.port, 2, (A, i, j))
    assert outer.tolist() == [0, 1, 2] and inner.tolist() == [0, 1] and vals.tolist() == [-5.0, 5.0]
    # M1 of the reference test: full 2x2 given row by row -> column-major storage
    A = np.array([1.0, 2.0, 3.0, 4.0]); i = np.array([0, 0, 1, 1], np.int32); j = np.array([0, 1, 0, 1], np.int32)
    outer, inner, vals = O.convert(O.port, 2, (A, i, j))
    assert outer.tolist() == [0, 2, 4] and inner.tolist() == [0, 1, 0, 1] and vals.tolist() == [1.0, 3.0, 2.0, 4.0]
    # out-of-order + duplicate mix
    A = np.array([1.0, 2.0, 2.5, 2.5, 4.0]); i = np.array([0, 1, 2, 2, 0], np.int32); j = np.array([0, 1, 2, 2, 2], np.int32)
    outer, inner, vals = O.convert(O.port, 3, (A, i, j))
    assert outer.tolist() == [0, 1, 2, 4]
    assert inner.tolist() == [0, 1, 0, 2]
    assert vals.tolist() == [1.0, 2.0, 4.0, 5.0]


def test_convert_empty():
    outer, inner, vals = O.convert(O.port, 4, (np.zeros(0), np.zeros(0, np.int32), np.zeros(0, np.int32)))
    assert outer.tolist() == [0] * 5 and len(inner) == 0


# --- literals from /root/reference/tests/test_time-integrator.cpp:29-65 (BDF-1 exact)
def test_bdf1_exact():
    n = 3
    dt = np.array([0.5, 0, 0, 0.0]); xk = np.array([1.0, 2.0, 4.0]); xh = np.zeros(4 * n)
    xh[:n] = [0.5, 1.0, 3.0]
    d = np.zeros(n)
    alpha = O.port.oracle_time_derivative(n, 1, dt, xk, xh, d)
    assert alpha == 2.0
    assert d.tolist() == [1.0, 2.0, 2.0]


# --- 3x3 KAT in the spirit of /root/reference/tests/test_linear-solver.cpp:26-61 (residual <= 1e-15)
def test_lu_3x3():
    A = np.array([[2.0, -1.0, 0.0], [-1.0, 2.0, -1.0], [0.0, -1.0, 2.0]])
    ii, jj = np.nonzero(A)
    csc = O.convert(O.port, 3, (A[ii, jj], ii.astype(np.int32), jj.astype(np.int32)))
    b = np.array([1.0, 0.0, 1.0]); x = np.zeros(3)
    assert O.port.oracle_lu_solve(3, *csc, b, x) == 0
    assert np.max(np.abs(A @ x - b)) <= 1e-15


@needs_ref
@pytest.mark.parametrize("seed", range(6))
def test_assembly_port_vs_reference_random(seed):
    n = [1, 7, 100, 1000, 5000, 20000][seed]
    J = P.random_triplets(n, 6 * n + 3, seed)
    M = P.random_triplets(n, n // 2 + 1, 100 + seed)
    alpha = 123.456 / (seed + 1)
    a = O.assemble_jb(O.port, n, J, M, alpha)
    b = O.assemble_jb(O.ref, n, J, M, alpha)
    for u, v in zip(a, b):
        assert u.tobytes() == v.tobytes()
    a = O.convert(O.port, n, J); b = O.convert(O.ref, n, J)
    for u, v in zip(a, b):
        assert u.tobytes() == v.tobytes()


@needs_ref
def test_assembly_edge_cases_vs_reference():
    n = 5
    empty = (np.zeros(0), np.zeros(0, np.int32), np.zeros(0, np.int32))
    J = (np.array([1.0, 0.0, -3.0, 3.0]), np.array([1, 4, 2, 2], np.int32), np.array([0, 4, 2, 2], np.int32))
    M = P.identity_mass(n)
    for jj, mm in [(J, M), (empty, M), (J, empty), (empty, empty), (M, M)]:
        a = O.assemble_jb(O.port, n, jj, mm, 7.0)
        b = O.assemble_jb(O.ref, n, jj, mm, 7.0)
        for u, v in zip(a, b):
            assert u.tobytes() == v.tobytes()


@needs_ref
@pytest.mark.parametrize("order", [1, 2, 3, 4])
def test_bdf_vs_reference(order):
    r = np.random.default_rng(order)
    n = 1000
    dt = r.uniform(0.01, 0.3, 4); xk = r.standard_normal(n); xh = r.standard_normal(4 * n)
    d1 = np.zeros(n); d2 = np.zeros(n)
    a1 = O.port.oracle_time_derivative(n, order, dt, xk, xh, d1)
    a2 = O.ref.daeref_time_derivative(n, order, dt, xk, xh, d2)
    assert a1 == a2
    assert d1.tobytes() == d2.tobytes()


@needs_ref
def test_residual_vs_reference():
    r = np.random.default_rng(5)
    n = 3000
    M = P.random_triplets(n, 4 * n, 11)
    dxdt = r.standard_normal(n); f = r.standard_normal(n)
    b_ref = np.zeros(n); b_port = np.zeros(n)
    O.ref.daeref_residual(n, len(M[0]), *[np.ascontiguousarray(a) for a in M], dxdt, f, b_ref)
    csc = O.convert(O.port, n, M)
    O.port.oracle_residual(n, *csc, dxdt, f, b_port)
    assert b_ref.tobytes() == b_port.tobytes()


@needs_ref
def test_lu_port_vs_reference_perovskite():
    N = 500
    x = P.perovskite_state(N)
    J = P.perovskite_jacobian(x, N); M = P.perovskite_mass(N)
    csc = O.assemble_jb(O.ref, 2 * N, J, M, 100.0)
    b = np.full(2 * N, 1.24)
    rc, x_ref, _, _ = O.ref_sparselu(2 * N, csc, b)
    assert rc == 0
    x_port = np.zeros(2 * N)
    assert O.port.oracle_lu_solve(2 * N, *csc, b, x_port) == 0
    eta_ref, _ = O.backward_error(2 * N, csc, x_ref, b)
    eta_port, _ = O.backward_error(2 * N, csc, x_port, b)
    assert eta_ref < 1e-14 and eta_port < 1e-14
    assert np.max(np.abs(x_port - x_ref)) <= 1e-8 * np.max(np.abs(x_ref))


def test_error_check_and_accept():
    n = 8
    dx = np.array([1e-9, -1e-9, 0, 0, 1e-3, 0, 0, 0.0]); x0 = np.ones(n); xk = np.ones(n)
    assert O.port.oracle_error_check_update(n, dx, x0, xk, 1e-6, 1e-6, 1e100) == 2
    assert xk[4] == 1.0 + 1e-3
    dx[:] = 1e-9; xk[:] = 1.0
    assert O.port.oracle_error_check_update(n, dx, x0, xk, 1e-6, 1e-6, 1e100) == 0
    dx[3] = np.nan; xk[:] = 1.0
    assert O.port.oracle_error_check_update(n, dx, x0, xk, 1e-6, 1e-6, 1e100) & 1
    assert xk[3] == 1.0 and xk[2] != 1.0  # early break at the offending element
    xh = np.zeros(4 * n); xh[:n] = 2.0; xk[:] = 1.0
    v = O.port.oracle_accept_step(n, xk, xh, 1, 1e-4)
    assert v == 0.5 and xh[:n].tolist() == [1.0] * n and xh[n:2 * n].tolist() == [2.0] * n
