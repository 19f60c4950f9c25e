"""GPU parity tests: the CUDA path behind the C ABI against the oracle on the same inputs.

Bars:
  assembly / residual / update / history : bit-exact (bytes compared)
  linear solve : normwise backward error <= 1e-12 and ||r||_2 <= 10 ||r_eigen||_2 + 1e-12 ||b||_2
"""
import numpy as np
import pytest

import oracle_lib as O
import problems as P
from gpu_lib import binding as B

pytestmark = pytest.mark.gpu

EMPTY = (np.zeros(0), np.zeros(0, np.int32), np.zeros(0, np.int32))


def same(a, b):
    return all(u.tobytes() == v.tobytes() for u, v in zip(a, b))


def checker():
    return O.ref if O.ref is not None else O.port


# ---------------------------------------------------------------- assembly (a3, a6)
@pytest.mark.parametrize("seed,n", [(0, 1), (1, 7), (2, 100), (3, 1000), (4, 20000), (5, 200000)])
def test_assembly_random_duplicates_bit_exact(seed, n):
    J = P.random_triplets(n, 6 * n + 3, seed)
    M = P.random_triplets(n, n // 2 + 1, 100 + seed)
    alpha = 123.456 / (seed + 1)
    c = B.Context(n)
    c.set_mass(M); c.set_jacobian(J)
    assert same(c.get_csc(0), O.convert(O.port, n, M))
    assert same(c.get_csc(1), O.convert(O.port, n, J))
    # a random sparsity pattern has no small separators (catastrophic fill for any direct
    # solver), so only the assembly stage of refactor runs here
    c.set_option("assemble_only", 1)
    c.refactor(alpha)
    assert same(c.get_csc(2), O.assemble_jb(checker(), n, J, M, alpha))
    c.close()


def test_assembly_edge_cases_bit_exact():
    n = 5
    J = (np.array([1.0, 0.0, -3.0, 3.0, -0.0]), np.array([1, 4, 2, 2, 3], np.int32), np.array([0, 4, 2, 2, 3], np.int32))
    M = P.identity_mass(n)
    for jj, mm in [(J, M), (EMPTY, M), (J, EMPTY), (M, M), (EMPTY, EMPTY)]:
        c = B.Context(n)
        c.set_mass(mm); c.set_jacobian(jj)
        c.refactor(7.0, allow=(B.EDECOMP,))
        assert same(c.get_csc(2), O.assemble_jb(checker(), n, jj, mm, 7.0))
        c.close()


def test_assembly_cached_pattern_hot_path():
    N = 400
    c = B.Context(2 * N)
    M = P.perovskite_mass(N)
    c.set_mass(M)
    for k in range(4):
        x = P.perovskite_state(N, seed=k)
        J = P.perovskite_jacobian(x, N)
        c.set_jacobian(J)
        c.refactor(10.0 * (k + 1))
        assert same(c.get_csc(2), O.assemble_jb(checker(), 2 * N, J, M, 10.0 * (k + 1)))
    assert c.stat("n_cold_assembly") == 2  # one for M, one for J; later calls reuse the pattern
    assert c.stat("n_symbolic") == 1
    # a different pattern of the same length must be detected and re-sorted
    A, i, j = P.perovskite_jacobian(x, N)
    i = i.copy(); j = j.copy(); i[5], j[5] = 7, 300
    c.set_jacobian((A, i, j))
    c.refactor(3.0, allow=(B.EDECOMP,))
    assert c.stat("n_cold_assembly") == 3 and c.stat("n_symbolic") == 2
    assert same(c.get_csc(2), O.assemble_jb(checker(), 2 * N, (A, i, j), M, 3.0))
    c.close()


def test_index_out_of_range_is_reported():
    c = B.Context(4)
    T = (np.ones(3), np.array([0, 4, 1], np.int32), np.array([0, 1, 1], np.int32))
    assert c.set_jacobian(T, allow=(B.EINDEX,)) == B.EINDEX
    T = (np.ones(3), np.array([0, 3, 1], np.int32), np.array([0, -1, 1], np.int32))
    assert c.set_mass(T, allow=(B.EINDEX,)) == B.EINDEX
    c.close()


# ---------------------------------------------------------------- residual (a4, a5)
@pytest.mark.parametrize("order", [1, 2, 3, 4])
def test_residual_bit_exact(order):
    r = np.random.default_rng(order)
    n = 5000
    M = P.random_triplets(n, 3 * n, 7)
    c = B.Context(n)
    c.set_mass(M)
    hist = [r.standard_normal(n) for _ in range(5)]  # xk, x[0..3]
    c.set_vector("xk", hist[0])
    for k in range(4):
        c.set_vector("x%d" % k, hist[k + 1])
    dt = r.uniform(0.01, 0.3, 4)
    f = r.standard_normal(n)
    alpha = c.residual(f, order, dt)
    xh = np.concatenate([hist[1], hist[2], hist[3], hist[4]])
    d = np.zeros(n); b = np.zeros(n)
    alpha_ref = O.port.oracle_time_derivative(n, order, dt, hist[0], xh, d)
    O.port.oracle_residual(n, *O.convert(O.port, n, M), d, f, b)
    assert alpha == alpha_ref
    assert c.get_vector("dxdt").tobytes() == d.tobytes()
    assert c.get_vector("b").tobytes() == b.tobytes()
    c.close()




# ---------------------------------------------------------------- linear solve (a7-a9)
def _solve_and_check(n, J, M, alpha, b=None, options=None):
    csc = O.assemble_jb(O.port, n, J, M, alpha)
    b = np.full(n, 1.24) if b is None else b
    c = B.Context(n)
    for k, v in (options or {}).items():
        c.set_option(k, v)
    c.set_mass(M); c.set_jacobian(J)
    c.refactor(alpha)
    c.set_rhs(b)
    x = c.solve()
    stats = {k: c.stat(k) for k in ("n_fronts", "n_levels", "nnz_factor", "max_front", "ms_last_refactor", "ms_last_solve",
                                    "n_units", "n_big_fronts", "n_launch_groups")}
    c.close()
    eta, r2 = O.backward_error(n, csc, x, b)
    assert eta <= 1e-12, (eta, stats)
    if O.ref is not None and n <= 300000:
        rc, x_ref, _, _ = O.ref_sparselu(n, csc, b)
        assert rc == 0
        _, r2_ref = O.backward_error(n, csc, x_ref, b)
        assert r2 <= 10.0 * r2_ref + 1e-12 * np.linalg.norm(b), (r2, r2_ref)
    return x, stats


def test_linear_solve_3x3_kat():
    # /root/reference/tests/test_linear-solver.cpp:26-61 style: per-entry residual <= 1e-15
    A = np.array([[2.0, -1.0, 0.0], [-1.0, 2.0, -1.0], [0.0, -1.0, 2.0]])
    ii, jj = np.nonzero(A)
    b = np.array([1.0, 0.0, 1.0])
    x, _ = _solve_and_check(3, (A[ii, jj], ii.astype(np.int32), jj.astype(np.int32)), EMPTY, 0.0, b)
    assert np.max(np.abs(A @ x - b)) <= 1e-15


def test_linear_solve_scalar_system():
    # C1: flame_propagation as shipped is N = 1
    x, _ = _solve_and_check(1, (np.array([-3.0]), np.zeros(1, np.int32), np.zeros(1, np.int32)),
                            P.identity_mass(1), 100.0, np.array([2.0]))
    assert x[0] == 2.0 / (-3.0 - 100.0)


@pytest.mark.parametrize("N,alpha", [(500, 100.0), (4000, 100.0), (4000, 1.0), (10000, 1.0)])
def test_linear_solve_perovskite(N, alpha):
    # N=10000 mirrors LinearSolver.BigSystem (/root/reference/tests/test_linear-solver.cpp:138-181):
    # b == 1.24 and per-entry |b - A x| <= 1e-7
    xs = P.perovskite_state(N)
    J = P.perovskite_jacobian(xs, N); M = P.perovskite_mass(N)
    x, stats = _solve_and_check(2 * N, J, M, alpha)
    csc = O.assemble_jb(O.port, 2 * N, J, M, alpha)
    _, r2 = O.backward_error(2 * N, csc, x, np.full(2 * N, 1.24))
    assert r2 / np.sqrt(2 * N) <= 1e-7


@pytest.mark.parametrize("n", [2, 3, 10, 1000, 100000])
def test_linear_solve_flame1d(n):
    u = P.flame1d_state(n)
    _solve_and_check(n, P.flame1d_jacobian(u), P.identity_mass(n), 50.0)


@pytest.mark.parametrize("m", [4, 33, 128])
def test_linear_solve_grid2d(m):
    _solve_and_check(m * m, P.grid2d_probe_matrix(m), EMPTY, 0.0)


@pytest.mark.parametrize("budget", [0, 2048, 16 * 1024, 96 * 1024])
@pytest.mark.parametrize("threads", [64, 256])
def test_linear_solve_schedule_variants(budget, threads):
    """Same answers whatever the partition into shared-memory units and per-level big fronts."""
    opts = {"unit_smem_bytes": budget, "unit_threads": threads, "chain": 0}  # the general multifrontal path
    N = 700
    xs = P.perovskite_state(N)
    _, st = _solve_and_check(2 * N, P.perovskite_jacobian(xs, N), P.perovskite_mass(N), 100.0, options=opts)
    assert (st["n_units"] == 0) == (budget == 0)
    _, st = _solve_and_check(48 * 48, P.grid2d_probe_matrix(48), EMPTY, 0.0, options=opts)
    if budget == 2048:
        assert st["n_big_fronts"] > 0 and st["n_units"] > 0  # mixed schedule
    n = 5000
    _solve_and_check(n, P.flame1d_jacobian(P.flame1d_state(n)), P.identity_mass(n), 50.0, options=opts)


def test_linear_solve_wide_groups():
    """serial_dim = 0 forces every front through the cooperative (8 / 32 / block wide) code."""
    N = 300
    xs = P.perovskite_state(N)
    _solve_and_check(2 * N, P.perovskite_jacobian(xs, N), P.perovskite_mass(N), 100.0, options={"serial_dim": 0, "chain": 0})
    _solve_and_check(40 * 40, P.grid2d_probe_matrix(40), EMPTY, 0.0, options={"serial_dim": 0, "unit_smem_bytes": 200 * 1024})


def test_repeated_refactor_and_solve_reuse_schedule():
    """Counters / tickets / epochs survive many launches on one context."""
    n = 20000
    u = P.flame1d_state(n)
    J = P.flame1d_jacobian(u); M = P.identity_mass(n)
    c = B.Context(n)
    c.set_option("chain", 0)  # the general path's tickets and epochs; the chain path has its own test file
    c.set_mass(M); c.set_jacobian(J)
    b = np.linspace(-1, 1, n)
    for k in range(12):
        alpha = 10.0 + k
        c.refactor(alpha)
        for _ in range(2):
            c.set_rhs(b)
            x = c.solve()
            eta, _ = O.backward_error(n, O.assemble_jb(O.port, n, J, M, alpha), x, b)
            assert eta <= 1e-12
    assert c.stat("n_symbolic") == 1
    c.close()


def test_linear_solve_dense_block_and_disconnected():
    # two disconnected dense 40x40 blocks plus 20 isolated unknowns
    r = np.random.default_rng(3)
    n = 100
    A = np.zeros((n, n))
    A[:40, :40] = r.standard_normal((40, 40)) + 40 * np.eye(40)
    A[40:80, 40:80] = r.standard_normal((40, 40)) + 40 * np.eye(40)
    A[80:, 80:] = np.diag(r.uniform(1, 2, 20))
    ii, jj = np.nonzero(A)
    _solve_and_check(n, (A[ii, jj], ii.astype(np.int32), jj.astype(np.int32)), EMPTY, 0.0, r.standard_normal(n))


def test_singular_matrix_reports_decomposition_failure():
    n = 4
    T = (np.array([1.0, 1.0, 1.0]), np.array([0, 1, 2], np.int32), np.array([0, 1, 2], np.int32))  # row/col 3 empty
    c = B.Context(n)
    c.set_mass(EMPTY); c.set_jacobian(T)
    assert c.refactor(0.0, allow=(B.EDECOMP,)) == B.EDECOMP
    c.close()


# ---------------------------------------------------------------- update / history (a11, a12)
def test_update_and_accept_bit_exact():
    r = np.random.default_rng(0)
    n = 100003
    x0 = r.standard_normal(n); x0[::7] = 0.0
    xk = x0 + 1e-3 * r.standard_normal(n)
    dx = 1e-7 * r.standard_normal(n) * 10.0 ** r.integers(-3, 3, n)
    # plant dx as the solution of I * dx = b
    c = B.Context(n)
    c.set_mass(EMPTY); c.set_jacobian(P.identity_mass(n)); c.refactor(0.0)
    for atol, rtol in [(1e-6, 1e-6), (1e-3, 1e-1), (1.0, 1e9)]:
        c.set_vector("xk", xk); c.set_vector("x0", x0); c.set_vector("b", dx)
        rc, flags = c.solve_update(atol, rtol)
        xk_ref = xk.copy()
        f_ref = O.port.oracle_error_check_update(n, dx, x0, xk_ref, atol, rtol, 1e100)
        assert rc == 0 and flags == f_ref
        assert c.get_xk().tobytes() == xk_ref.tobytes()
    dx2 = dx.copy(); dx2[n // 2] = np.nan
    c.set_vector("xk", xk); c.set_vector("b", dx2)
    rc, flags = c.solve_update(1e-6, 1e-6)
    assert flags & B.DIVERGED
    dx2[n // 2] = 1e101
    c.set_vector("xk", xk); c.set_vector("b", dx2)
    assert c.solve_update(1e-6, 1e-6)[1] & B.DIVERGED
    # rollback restores xk = x[0]
    c.rollback()
    assert c.get_xk().tobytes() == x0.tobytes()
    # history shift + variability
    hist = [r.standard_normal(n) for _ in range(4)]
    for k in range(4):
        c.set_vector("x%d" % k, hist[k])
    c.set_vector("xk", xk)
    v = c.accept_step(True, 1e-4)
    xh = np.concatenate(hist)
    v_ref = O.port.oracle_accept_step(n, xk, xh, 1, 1e-4)
    assert v == v_ref
    for k in range(4):
        assert c.get_vector("x%d" % k).tobytes() == xh[k * n:(k + 1) * n].tobytes()
    assert c.accept_step(False, 1e-4) == 0.0
    c.close()


# ---------------------------------------------------------------- one BDF step end to end
def test_newton_iterations_perovskite_vs_oracle():
    """Three quasi-Newton iterations of the first BDF-1 step of C2 (N reduced): the device
    pipeline (set_jacobian -> residual -> refactor -> solve_update) against the same sequence
    done with the oracle; b is bit-exact, dx within the linear-solve bar."""
    N = 300; n = 2 * N
    x0 = np.concatenate([np.ones(N), np.zeros(N)])
    M = P.perovskite_mass(N)
    dt = np.array([0.01, 0, 0, 0.0]); t = 0.01

    def rhs(x):
        h = 1.0 / (N - 1); invh2 = 1.0 / (h * h)
        f = np.zeros(n); Pc = x[:N]; Ph = x[N:]
        f[1:N - 1] = (Pc[2:] - 2 * Pc[1:-1] + Pc[:-2] + 0.5 * ((Pc[2:] + Pc[1:-1]) * (Ph[2:] - Ph[1:-1]) -
                      (Pc[1:-1] + Pc[:-2]) * (Ph[1:-1] - Ph[:-2]))) * invh2
        f[0] = (Pc[1] - Pc[0] + 0.5 * (Pc[1] + Pc[0]) * (Ph[1] - Ph[0])) * invh2
        f[N - 1] = -(Pc[-1] - Pc[-2] + 0.5 * (Pc[-1] + Pc[-2]) * (Ph[-1] - Ph[-2])) * invh2
        f[N + 1:n - 1] = (Ph[2:] - 2 * Ph[1:-1] + Ph[:-2]) * invh2 - (1.0 - Pc[1:-1])
        f[N] = Ph[0] + t; f[n - 1] = Ph[-1] - t
        return f

    c = B.Context(n)
    c.set_state(x0); c.set_mass(M)
    xk_ref = x0.copy(); xh = np.concatenate([x0, np.zeros(3 * n)])
    mcsc = O.convert(O.port, n, M)
    for it in range(3):
        xk = c.get_xk()
        assert np.max(np.abs(xk - xk_ref)) <= 1e-9 * max(1.0, np.max(np.abs(xk_ref)))
        f = rhs(xk)
        alpha = c.residual(f, 1, dt)
        d = np.zeros(n); b_ref = np.zeros(n)
        assert alpha == O.port.oracle_time_derivative(n, 1, dt, xk, xh, d)
        O.port.oracle_residual(n, *mcsc, d, f, b_ref)
        assert c.get_vector("b").tobytes() == b_ref.tobytes()
        J = P.perovskite_jacobian(xk, N)
        c.set_jacobian(J); c.refactor(alpha)
        csc = O.assemble_jb(O.port, n, J, M, alpha)
        assert same(c.get_csc(2), csc)
        rc, flags = c.solve_update(1e-6, 1e-6)
        dx = c.get_vector("dx")
        eta, _ = O.backward_error(n, csc, dx, b_ref)
        assert rc == 0 and eta <= 1e-12
        dx_ref = np.zeros(n)
        assert O.port.oracle_lu_solve(n, *csc, b_ref, dx_ref) == 0
        assert np.max(np.abs(dx - dx_ref)) <= 1e-8 * max(np.max(np.abs(dx_ref)), 1e-300)
        xk_ref = xk.copy()
        O.port.oracle_error_check_update(n, dx, x0, xk_ref, 1e-6, 1e-6, 1e100)
    c.close()


def test_two_live_contexts_with_different_unit_sizes():
    """The shared-memory opt-in of the unit kernels belongs to the function, not to a context (subtree.cu: unit_setup):
    a context with small units set up while one with big units is alive must not shrink it under the other's feet."""
    m1, m2 = 128, 64
    J1, J2 = P.grid2d_probe_matrix(m1), P.grid2d_probe_matrix(m2)
    c1 = B.Context(m1 * m1)
    c1.set_mass(EMPTY); c1.set_jacobian(J1); c1.refactor(0.0)
    c2 = B.Context(m2 * m2)
    c2.set_option("unit_smem_bytes", 4096)
    c2.set_mass(EMPTY); c2.set_jacobian(J2); c2.refactor(0.0)
    c1.refactor(0.0)  # used to fail: invalid argument (dynamic shared memory above the opt-in the second context left)
    b = np.full(m1 * m1, 1.24)
    c1.set_rhs(b)
    x = c1.solve()
    eta, _ = O.backward_error(m1 * m1, O.assemble_jb(O.port, m1 * m1, J1, EMPTY, 0.0), x, b)
    assert eta <= 1e-12
    c1.close(); c2.close()


@pytest.mark.parametrize("m,budget", [(33, 2048), (48, 2048), (64, 4096), (128, -1), (200, -1)])
def test_linear_solve_blocked_big_fronts(m, budget):
    """Big fronts through the blocked multi-CTA LU (panel / trsm / gemm kernels of numeric.cu): pivot
    blocks that are not multiples of the panel width, tiles that straddle the L panel / U panel /
    contribution block boundary, many fronts per level."""
    opts = {"blocked_lu": 1, "blocked_lu_min_dim": 8}
    if budget >= 0:
        opts["unit_smem_bytes"] = budget
    x1, st = _solve_and_check(m * m, P.grid2d_probe_matrix(m), EMPTY, 0.0, options=opts)
    assert st["n_big_fronts"] > 0
    opts["blocked_lu"] = 0
    x0, _ = _solve_and_check(m * m, P.grid2d_probe_matrix(m), EMPTY, 0.0, options=opts)
    assert np.max(np.abs(x1 - x0)) <= 1e-10 * np.max(np.abs(x0))


@pytest.mark.parametrize("n", [1, 3, 255, 6149])
def test_residual_diagonal_mass_odd_sizes(n):
    """k_residual_diag works on 16-byte pairs: vectors of odd length must still start aligned."""
    rng = np.random.default_rng(n)
    x = rng.standard_normal(n); f = rng.standard_normal(n)
    M = (rng.uniform(0.5, 2.0, n), np.arange(n, dtype=np.int32), np.arange(n, dtype=np.int32))
    c = B.Context(n)
    c.set_state(x); c.set_mass(M)
    c.set_vector("xk", x + 1e-3 * rng.standard_normal(n))
    dt = np.array([1e-2, 0.0, 0.0, 0.0])
    alpha = c.residual(f, 1, dt)
    xk = c.get_vector("xk"); xh = np.concatenate([x, np.zeros(3 * n)])
    d = np.zeros(n); b_ref = np.zeros(n)
    assert alpha == O.port.oracle_time_derivative(n, 1, dt, xk, xh, d)
    O.port.oracle_residual(n, *O.convert(O.port, n, M), d, f, b_ref)
    assert c.get_vector("b").tobytes() == b_ref.tobytes()
    c.close()


def test_big_front_fused_panel_is_bit_identical():
    """One-launch panel (every CTA factors the diagonal block for itself) vs diag / rows / trsm launches."""
    m = 160
    n = m * m
    u = P.grid2d_state(m)
    J = P.grid2d_jacobian(u, m)
    M = P.identity_mass(n)
    b = np.full(n, 1.24)
    out = []
    for fused in (1, 0):
        c = B.Context(n)
        c.set_option("chain", 0); c.set_option("bf_fused_panel", fused)
        c.set_mass(M); c.set_jacobian(J)
        c.refactor(50.0)
        c.set_rhs(b)
        out.append(c.solve().tobytes())
        assert c.stat("max_front") >= 96
        c.close()
    assert out[0] == out[1]


@pytest.mark.parametrize("m", [160, 300])
def test_big_front_single_pass_contribution_block_is_bit_identical(m):
    """The contribution block of a big front updated once behind the last block column (k_bf_schur: block columns applied in
    order, each as old - sum of its 32 products) vs once per block column by the trailing update: the same bits."""
    n = m * m
    u = P.grid2d_state(m)
    J = P.grid2d_jacobian(u, m)
    M = P.identity_mass(n)
    b = np.full(n, 1.24)
    out = []
    for schur in (1, 0):
        c = B.Context(n)
        c.set_option("chain", 0); c.set_option("bf_schur", schur)
        c.set_mass(M); c.set_jacobian(J)
        c.refactor(50.0)
        c.set_rhs(b)
        out.append((c.solve().tobytes(), c.stat("kernels_launched")))
        assert c.stat("max_front") >= 96
        c.close()
    assert out[0][0] == out[1][0]
    assert out[0][1] > out[1][1]  # (one more launch per level of big fronts: the single pass really ran)


def test_big_front_dmma_matches_fma_tiles():
    """Trailing update on the FP64 tensor cores (mma.m8n8k4.f64) vs the FP64 FMA tiles: same factorisation up to
    the summation order inside a 32-deep dot product; both must meet the backward-error bar of the linear solve."""
    m = 160
    n = m * m
    u = P.grid2d_state(m)
    J = P.grid2d_jacobian(u, m)
    M = P.identity_mass(n)
    b = np.full(n, 1.24)
    csc = O.assemble_jb(O.port, n, J, M, 50.0)
    out = []
    for dmma in (1, 0):
        c = B.Context(n)
        c.set_option("chain", 0); c.set_option("bf_dmma", dmma)
        c.set_mass(M); c.set_jacobian(J)
        c.refactor(50.0)
        c.set_rhs(b)
        x = c.solve()
        eta, _ = O.backward_error(n, csc, x, b)
        assert eta <= 1e-12, (dmma, eta)
        assert c.stat("max_front") >= 96
        out.append(x)
        c.close()
    assert np.max(np.abs(out[0] - out[1])) <= 1e-10 * np.max(np.abs(out[1]))


@pytest.mark.parametrize("m", [160, 512])
def test_linear_solve_grid2d_default_options(m):
    """C4-style Jacobian through the default path (unit kernels below, blocked big fronts above).  Levels mix
    fronts with different pivot counts: a front with fewer pivots than the level's maximum has the larger
    trailing matrix in its last block column, which the launch grids must cover (m = 160 caught them short)."""
    n = m * m
    u = P.grid2d_state(m)
    x, st = _solve_and_check(n, P.grid2d_jacobian(u, m), P.identity_mass(n), 50.0, options={"chain": 0})
    assert st["n_big_fronts"] > 0


def _grid_solve_bytes(m, options, zero_row=None):
    n = m * m
    u = P.grid2d_state(m)
    A, I, Jc = P.grid2d_jacobian(u, m)
    if zero_row is not None:
        A = A.copy(); A[I == zero_row] = 0.0
    c = B.Context(n)
    c.set_option("chain", 0)
    for k, v in options.items():
        c.set_option(k, v)
    c.set_mass(P.identity_mass(n) if zero_row is None else EMPTY); c.set_jacobian((A, I, Jc))
    rc = c.refactor(50.0 if zero_row is None else 0.0, allow=(B.EDECOMP,))
    out = None
    if rc == 0:
        c.set_rhs(np.full(n, 1.24))
        out = c.solve().tobytes()
    assert c.stat("max_front") >= 96
    c.close()
    return rc, out


@pytest.mark.parametrize("m", [160, 300])
def test_big_front_latency_variants_are_bit_identical(m):
    """Round-2 latency work on the big fronts -- triangular sweeps with the pivot unknowns in shared memory and batched
    loads (bf_fast_solve), the factored diagonal block handed from the panel launch to the trailing update instead of
    being factored twice (bf_diag_copy), the next pivot's reciprocal formed beside the update (bf_lu_lookahead) -- performs
    the same operations in the same order: each switch alone and all together give the bits of the round-1 kernels."""
    base = {"bf_fast_solve": 0, "bf_diag_copy": 0, "bf_lu_lookahead": 0}
    rc0, ref = _grid_solve_bytes(m, base)
    assert rc0 == 0
    for on in (("bf_fast_solve",), ("bf_diag_copy",), ("bf_lu_lookahead",), ("bf_fast_solve", "bf_diag_copy", "bf_lu_lookahead")):
        rc, got = _grid_solve_bytes(m, {**base, **{k: 1 for k in on}})
        assert rc == 0 and got == ref, on
    rc, got = _grid_solve_bytes(m, {**base, "bf_diag_copy": 1, "bf_lu_lookahead": 1, "bf_fused_panel": 0})
    assert rc == 0 and got == ref


@pytest.mark.parametrize("max_super", [100, 512])
def test_big_front_latency_variants_other_pivot_block_sizes(max_super):
    """Pivot blocks that are not multiples of the 32-pivot step (max_super = 100) and fronts with more than eight steps
    (max_super = 512: a warp of the triangular sweeps reloads its registers for the block eight steps ahead)."""
    base = {"bf_fast_solve": 0, "bf_diag_copy": 0, "bf_lu_lookahead": 0, "max_super": max_super}
    rc0, ref = _grid_solve_bytes(400, base)
    rc, got = _grid_solve_bytes(400, {"max_super": max_super})
    assert rc0 == 0 and rc == 0 and got == ref


@pytest.mark.parametrize("opts", [{}, {"bf_diag_copy": 0, "bf_lu_lookahead": 0}, {"bf_diag_copy": 1, "bf_lu_lookahead": 0},
                                  {"bf_diag_copy": 0, "bf_lu_lookahead": 1}, {"bf_fused_panel": 0}])
def test_big_front_zero_pivot_is_reported(opts):
    """A numerically empty row in the top separator: its pivot is exactly zero when its block column is reached, whichever
    launch factors the diagonal block (panel CTA that keeps the copy, trailing update's extra CTA, separate launch)."""
    m = 160
    mid = (m // 2) * m + m // 2
    rc, _ = _grid_solve_bytes(m, opts, zero_row=mid)
    assert rc == B.EDECOMP
    rc, _ = _grid_solve_bytes(m, opts, zero_row=m * m - 1)
    assert rc == B.EDECOMP


def test_big_front_side_stream_variants_are_bit_identical():
    """The side stream of a context: big-front storage prepared under the unit launches (bf_zero_ahead) and unit launches
    that share a stage of the tree with a level of big fronts run beside it (bf_overlap_units).  Scheduling only: the same
    bits with either switch off, and from one run to the next."""
    m = 400
    rc0, ref = _grid_solve_bytes(m, {"bf_zero_ahead": 0, "bf_overlap_units": 0})
    assert rc0 == 0
    for opts in ({}, {}, {"bf_zero_ahead": 0}, {"bf_overlap_units": 0}, {}):
        rc, got = _grid_solve_bytes(m, opts)
        assert rc == 0 and got == ref, opts
