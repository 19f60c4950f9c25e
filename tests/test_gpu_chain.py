"""GPU parity tests of the block-chain path (dae-cpp_b200/csrc/chain.cu) through the C ABI: banded
patterns (flame-1D, perovskite) are factored and solved with no per-front tables.  Same bars as
test_gpu_parity.py: assembly bit-exact, linear solve eta <= 1e-12 and residual within 10x of
Eigen::SparseLU's, update loop bit-exact given dx."""
import numpy as np
import pytest

import oracle_lib as O
import problems as P
from gpu_lib import binding as B

pytestmark = pytest.mark.gpu

EMPTY = (np.zeros(0), np.zeros(0, np.int32), np.zeros(0, np.int32))


def solve_chain(n, J, M, alpha, b=None, expect_block=None, options=None):
    csc = O.assemble_jb(O.port, n, J, M, alpha)
    b = np.full(n, 1.24) if b is None else b
    c = B.Context(n)
    c.set_option("chain_min_n", 1)
    for k, v in (options or {}).items():
        c.set_option(k, v)
    c.set_mass(M); c.set_jacobian(J)
    c.refactor(alpha)
    got = c.get_csc(2)
    assert all(u.tobytes() == v.tobytes() for u, v in zip(got, csc)), "assembly not bit-exact"
    c.set_rhs(b)
    x = c.solve()
    blk = int(c.stat("chain_block"))
    stats = {k: c.stat(k) for k in ("chain_block", "chain_levels", "chain_tiles", "chain_identity_perm", "nnz_factor")}
    c.close()
    if expect_block is not None:
        assert blk == expect_block, stats
    eta, r2 = O.backward_error(n, csc, x, b)
    assert eta <= 1e-12, (eta, stats)
    if O.ref is not None and n <= 300000:
        rc, x_ref, _, _ = O.ref_sparselu(n, csc, b)
        assert rc == 0
        _, r2_ref = O.backward_error(n, csc, x_ref, b)
        assert r2 <= 10.0 * r2_ref + 1e-12 * np.linalg.norm(b), (r2, r2_ref)
    return x, stats


@pytest.mark.parametrize("n", [2, 3, 17, 1000, 2048, 2049, 100000, 1000000])
def test_chain_flame1d(n):
    u = P.flame1d_state(n)
    _, st = solve_chain(n, P.flame1d_jacobian(u), P.identity_mass(n), 50.0, expect_block=1)
    assert st["chain_identity_perm"] == 1
    assert st["chain_levels"] == (1 if n <= 2048 else 2)


def test_chain_three_grid_levels():
    n = 2048 * 2048 + 77
    u = P.flame1d_state(n)
    _, st = solve_chain(n, P.flame1d_jacobian(u), P.identity_mass(n), 50.0, expect_block=1)
    assert st["chain_levels"] == 3


@pytest.mark.parametrize("N,alpha", [(5, 100.0), (300, 100.0), (4000, 100.0), (4000, 1.0), (10000, 1.0)])
def test_chain_perovskite(N, alpha):
    xs = P.perovskite_state(N)
    _, st = solve_chain(2 * N, P.perovskite_jacobian(xs, N), P.perovskite_mass(N), alpha, expect_block=2)
    assert st["chain_identity_perm"] == 0


def test_chain_block3_random_band_with_duplicates():
    n = 3001
    rng = np.random.default_rng(5)
    I, Jc, A = [], [], []
    for i in range(n):
        for j in range(max(0, i - 4), min(n, i + 5)):
            if abs(i // 3 - j // 3) <= 1 and rng.random() < 0.6:
                for _ in range(1 + (rng.random() < 0.3)):  # duplicates are summed in input order
                    I.append(i); Jc.append(j); A.append(rng.standard_normal())
        I.append(i); Jc.append(i); A.append(9.0 + rng.random())
    p = rng.permutation(len(A))
    J = (np.array(A)[p], np.array(I, np.int32)[p], np.array(Jc, np.int32)[p])
    solve_chain(n, J, P.identity_mass(n), 2.0, b=rng.standard_normal(n), expect_block=3)


def test_chain_matches_general_path():
    N = 700
    xs = P.perovskite_state(N)
    J = P.perovskite_jacobian(xs, N); M = P.perovskite_mass(N)
    b = np.random.default_rng(0).standard_normal(2 * N)
    x1, _ = solve_chain(2 * N, J, M, 10.0, b=b, expect_block=2)
    x0, _ = solve_chain(2 * N, J, M, 10.0, b=b, expect_block=0, options={"chain": 0})
    assert np.max(np.abs(x1 - x0)) <= 1e-9 * np.max(np.abs(x0))


def test_chain_singular_matrix_reports_decomposition_failure():
    n = 1000
    A = np.ones(3 * n - 2)
    i = np.concatenate([np.arange(n), np.arange(1, n), np.arange(n - 1)]).astype(np.int32)
    j = np.concatenate([np.arange(n), np.arange(n - 1), np.arange(1, n)]).astype(np.int32)
    A[:n] = 2.0
    A[500] = 0.0; A[n + 499] = 0.0; A[2 * n - 1 + 500] = 0.0  # row 500 is identically zero (entries kept in the pattern)
    c = B.Context(n)
    c.set_option("chain_min_n", 1)
    c.set_mass(EMPTY); c.set_jacobian((A, i, j))
    rc = c.refactor(0.0, allow=(B.EDECOMP,))
    assert int(c.stat("chain_block")) == 1
    assert rc == B.EDECOMP
    c.close()


def test_chain_newton_iterations_match_oracle():
    # whole resident iterations on the chain path, checked like test_newton_iterations_perovskite_vs_oracle
    N = 1500
    n = 2 * N
    M = P.perovskite_mass(N)
    x = P.perovskite_state(N)
    c = B.Context(n)
    c.set_state(x)
    c.set_mass(M)
    dt4 = np.array([1e-3, 0.0, 0.0, 0.0])
    for it in range(3):
        xk = c.get_xk()
        J = P.perovskite_jacobian(xk, N)
        c.set_jacobian(J)
        f = np.sin(np.arange(n) * 0.01) * 1e-3
        alpha = c.residual(f, 1, dt4)
        xk_before = c.get_vector("xk")
        rc, flags = c.iteration_resident(1, 1, dt4, 1e-6, 1e-6)
        assert rc == 0
        assert int(c.stat("chain_block")) == 2
        csc = O.assemble_jb(O.port, n, J, M, alpha)
        dx = c.get_vector("dx"); b = c.get_vector("b")
        eta, _ = O.backward_error(n, csc, dx, b)
        assert eta <= 1e-12
        assert (xk_before + dx).tobytes() == c.get_vector("xk").tobytes()
        want = 0
        e = np.abs(dx)
        x0 = np.abs(c.get_vector("x0"))
        if np.any(e > 1e100) or np.any(np.isnan(dx)): want |= 1
        if np.any((x0 > 1e-14) & (e / np.where(x0 > 0, x0, 1.0) > 1e-6)) or np.any(e > 1e-6): want |= 2
        assert flags == want
    c.close()


@pytest.mark.parametrize("fuse", [0, 1, 2, 3])
def test_chain_top_level_fused_into_last_cta(fuse):
    # two levels: the top level rides in the last CTA of the level-0 launches (fuse) or has its own launches
    n = 300000
    u = P.flame1d_state(n)
    x, st = solve_chain(n, P.flame1d_jacobian(u), P.identity_mass(n), 50.0, expect_block=1, options={"chain_fuse_top": fuse})
    assert st["chain_levels"] == 2
    N = 6000
    xs = P.perovskite_state(N)
    solve_chain(2 * N, P.perovskite_jacobian(xs, N), P.perovskite_mass(N), 10.0, expect_block=2, options={"chain_fuse_top": fuse})


def test_chain_repeated_iterations_reuse_counters():
    n = 50000
    u = P.flame1d_state(n)
    J = P.flame1d_jacobian(u); M = P.identity_mass(n)
    c = B.Context(n)
    c.set_mass(M); c.set_jacobian(J)
    b = np.linspace(-1, 1, n)
    for k in range(6):
        alpha = 10.0 + k
        c.refactor(alpha)
        csc = O.assemble_jb(O.port, n, J, M, alpha)
        for _ in range(2):
            c.set_rhs(b)
            x = c.solve()
            eta, _ = O.backward_error(n, csc, x, b)
            assert eta <= 1e-12
    c.close()


# (options, top level solved by cyclic reduction?)  Within each group only the scheduling differs.
VARIANTS = [
    ({}, True),  # defaults: edge tiles affine, packed staging, cyclic-reduction top level, fused residual, lean flags
    ({"chain_fuse_residual": 0}, True),
    ({"chain_fuse_residual": 1, "keep_b": 1}, True),
    ({"lean_flags": 0}, True),
    ({"chain_packed": 0}, True),
    ({"chain_affine": 0}, True),
    ({"chain_top_pcr": 0}, False),                        # top level by LU in an extra CTA of the forward launch
    ({"chain_top_pcr": 0, "chain_defer_top": 0}, False),  # top level by LU in its own launch
    ({"chain_fuse_top": 0}, False),
    ({"chain_fuse_top": 3}, False),
    ({"chain_top_pcr": 0, "chain_defer_top": 0, "chain_fuse_residual": 0, "lean_flags": 0, "chain_packed": 0, "chain_affine": 0,
      "chain_fuse_top": 0}, False),
]


def _fused_iterations(problem, n, options, iters=3):
    """Three fused Newton iterations (daecpp_b200_iteration_resident) and everything they leave behind."""
    if problem == "flame1d":
        x = P.flame1d_state(n); M = P.identity_mass(n); J = P.flame1d_jacobian(x); f = P.flame1d_rhs(x)
    else:
        N = n // 2
        x = P.perovskite_state(N); M = P.perovskite_mass(N); J = P.perovskite_jacobian(x, N)
        f = np.sin(np.arange(n) * 0.01) * 1e-3
    c = B.Context(n)
    for k, v in options.items():
        c.set_option(k, v)
    c.set_state(x); c.set_mass(M); c.set_jacobian(J)
    dt4 = np.array([1e-3, 0.0, 0.0, 0.0])
    c.residual(f, 1, dt4)  # uploads f (and leaves b behind for the variants that do not rewrite it)
    out = []
    for it in range(iters):
        rc, flags = c.iteration_resident(0 if it == 1 else 1, 1, dt4, 1e-6, 1e-6)  # iteration 1 reuses the factors
        out.append((rc, flags, c.get_vector("dx"), c.get_vector("xk")))
    b = c.get_vector("b").tobytes()
    c.close()
    return out, b


@pytest.mark.parametrize("problem,n", [("flame1d", 50000), ("flame1d", 2048 * 3 + 5), ("flame1d", 1000000), ("perovskite", 8000)])
def test_fused_iteration_variants_agree(problem, n):
    """Staging variants, where the top level is factored, the fused residual and the flag delivery change
    how the iteration is scheduled, never what it computes: bit-identical.  The cyclic-reduction top level
    is different arithmetic on the same reduced system: it agrees with the LU top level to rounding."""
    ref = {}
    for opt, pcr in VARIANTS:
        if n == 1000000 and len(opt) > 1:
            continue
        got, _ = _fused_iterations(problem, n, opt)
        assert all(r[0] == 0 for r in got), opt
        if pcr not in ref:
            ref[pcr] = got
            continue
        for it, (a, g) in enumerate(zip(ref[pcr], got)):
            assert a[1] == g[1], (opt, it, a[1], g[1])
            assert a[2].tobytes() == g[2].tobytes(), (opt, it, "dx differs")
            assert a[3].tobytes() == g[3].tobytes(), (opt, it, "xk differs")
    for it, (a, g) in enumerate(zip(ref[True], ref[False])):
        scale = np.max(np.abs(g[2]))
        assert np.max(np.abs(a[2] - g[2])) <= 1e-10 * scale, (it, np.max(np.abs(a[2] - g[2])), scale)
        assert a[1] == g[1]


def test_fused_residual_keep_b_matches_residual_kernel():
    n = 30000
    _, b_kernel = _fused_iterations("flame1d", n, {"chain_fuse_residual": 0}, iters=2)
    _, b_fused = _fused_iterations("flame1d", n, {"chain_fuse_residual": 1, "keep_b": 1}, iters=2)
    assert b_kernel == b_fused


def test_chain_trace_layout():
    n = 2048 * 5
    x = P.flame1d_state(n)
    c = B.Context(n)
    c.set_state(x); c.set_mass(P.identity_mass(n)); c.set_jacobian(P.flame1d_jacobian(x))
    dt4 = np.array([1e-3, 0.0, 0.0, 0.0])
    c.residual(P.flame1d_rhs(x), 1, dt4)
    c.iteration_resident(1, 1, dt4, 1e-6, 1e-6)
    c.set_option("trace_units", 1)
    c.iteration_resident(1, 1, dt4, 1e-6, 1e-6)
    tiles = int(c.stat("chain_tiles"))
    tr = c.unit_trace().ravel().reshape(5, tiles, 8)
    c.close()
    assert tiles == 5
    assert np.all(tr[0, :, 0] > 0) and np.all(tr[0, :, 4] >= tr[0, :, 0])      # factor tiles: start <= done
    assert np.all(tr[2, :, 0] >= tr[0, :, 4].min()) and np.all(tr[4, :, 3] >= tr[2, :, 4].max())  # factor -> forward -> backward


def test_iterations_resident_equals_single_iterations():
    n = 40000
    x = P.flame1d_state(n)
    dt4 = np.array([1e-3, 0.0, 0.0, 0.0])
    res = []
    for batched in (False, True):
        for spin in (1, 0):
            c = B.Context(n)
            c.set_option("spin_flags", spin)
            c.set_state(x); c.set_mass(P.identity_mass(n)); c.set_jacobian(P.flame1d_jacobian(x))
            c.residual(P.flame1d_rhs(x), 1, dt4)
            if batched:
                rc, flags = c.iterations_resident(4, 1, 1, dt4, 1e-6, 1e-6)
            else:
                for _ in range(4):
                    rc, flags = c.iteration_resident(1, 1, dt4, 1e-6, 1e-6)
            res.append((rc, flags, c.get_vector("xk").tobytes(), c.get_vector("dx").tobytes()))
            c.close()
    assert all(r == res[0] for r in res[1:])


@pytest.mark.parametrize("problem,n,count", [("flame1d", 200000, 9), ("perovskite", 8000, 6), ("flame1d", 3000, 2)])
def test_pipelined_resident_iterations_equal_unpipelined(problem, n, count):
    """daecpp_b200_iterations_resident enqueues iteration k+1 behind k before the host has read k's flags (api.cu): the
    iterates, updates, flags and the refactorisation count must be those of the one-at-a-time loop, bit for bit."""
    if problem == "flame1d":
        x = P.flame1d_state(n); M = P.identity_mass(n); J = P.flame1d_jacobian(x); f = P.flame1d_rhs(x)
    else:
        N = n // 2
        x = P.perovskite_state(N); M = P.perovskite_mass(N); J = P.perovskite_jacobian(x, N)
        f = np.sin(np.arange(n) * 0.01) * 1e-3
    dt4 = np.array([1e-3, 0.0, 0.0, 0.0])
    res = []
    for pipe in (1, 0):
        for refactor in (1, 0):
            c = B.Context(n)
            c.set_option("pipeline_iterations", pipe)
            c.set_state(x); c.set_mass(M); c.set_jacobian(J)
            c.residual(f, 1, dt4)
            rc0, _ = c.iteration_resident(1, 1, dt4, 1e-6, 1e-6)  # factors for the refactor = 0 variant
            rc, flags = c.iterations_resident(count, refactor, 1, dt4, 1e-6, 1e-6)
            res.append((refactor, rc0, rc, flags, c.stat("n_refactor"), c.get_vector("xk").tobytes(), c.get_vector("dx").tobytes()))
            c.close()
    assert res[0] == res[2] and res[1] == res[3]
    assert res[0][4] == count + 1 and res[1][4] == 1


def test_pipelined_resident_iterations_stop_where_the_host_must_act():
    """A factorisation that fails inside a pipelined block: the iteration queued behind it is cancelled on the device, the
    call returns EDECOMP and leaves the iterate exactly where the one-at-a-time loop leaves it."""
    n = 50000
    x = P.flame1d_state(n)
    A, i, j = P.flame1d_jacobian(x)
    M = P.identity_mass(n)
    dt4 = np.array([1e-3, 0.0, 0.0, 0.0])
    res = []
    for pipe in (1, 0):
        c = B.Context(n)
        c.set_option("pipeline_iterations", pipe)
        c.set_state(x); c.set_mass(M); c.set_jacobian((A, i, j))
        c.residual(P.flame1d_rhs(x), 1, dt4)
        rc, flags = c.iterations_resident(3, 1, 1, dt4, 1e-6, 1e-6)
        assert rc == 0
        bad = A.copy()
        bad[(i == j)] = 1.0 / dt4[0]  # J_ii = alpha * M_ii: the diagonal of Jb = J - alpha*M cancels, the off-diagonals stay
        bad[(i != j)] = 0.0           # ... and with them gone Jb = 0
        c.set_jacobian_values_raw(np.ascontiguousarray(bad), len(bad))
        rc, flags = c.iterations_resident(5, 1, 1, dt4, 1e-6, 1e-6, allow=(B.EDECOMP,))
        res.append((rc, c.stat("n_refactor"), c.get_vector("xk").tobytes()))
        c.set_jacobian_values_raw(np.ascontiguousarray(A), len(A))      # and the context is usable again afterwards
        rc2, _ = c.iterations_resident(2, 1, 1, dt4, 1e-6, 1e-6)
        res[-1] += (rc2, c.get_vector("xk").tobytes())
        c.close()
    assert res[0][0] == B.EDECOMP
    assert res[0] == res[1]


@pytest.mark.parametrize("problem,n", [("flame1d", 30001), ("perovskite", 8000)])
def test_iteration_host_direct_xk_store(problem, n):
    """Small systems: xk reaches the caller's pinned buffer by stores of the backward kernel (no copy node,
    no stream synchronisation) -- same bits as the device vector and as the copy path."""
    if problem == "flame1d":
        x = P.flame1d_state(n); M = P.identity_mass(n); J = P.flame1d_jacobian(x); f = P.flame1d_rhs(x)
    else:
        N = n // 2
        x = P.perovskite_state(N); M = P.perovskite_mass(N); J = P.perovskite_jacobian(x, N)
        f = np.sin(np.arange(n) * 0.01) * 1e-3
    lib = B.load()
    dt4 = np.array([1e-3, 0.0, 0.0, 0.0])
    res = []
    for direct in (1, 0):
        c = B.Context(n, 0, lib)
        c.set_option("pinned_io", 1); c.set_option("host_xk_direct", direct)
        pf = B.PinnedArray(lib, (n,), np.float64); px = B.PinnedArray(lib, (n,), np.float64)
        pf.array[:] = f; px.array[:] = -7.0
        c.set_state(x); c.set_mass(M); c.set_jacobian(J)
        hist = []
        for it in range(3):
            rc, flags, alpha = c.iteration_host(pf.array, 0 if it == 1 else 1, 1, dt4, 1e-6, 1e-6, px.array)
            assert rc == 0
            assert px.array.tobytes() == c.get_vector("xk").tobytes()
            hist.append((flags, alpha, px.array.tobytes()))
        res.append(hist)
        c.close(); pf.free(); px.free()
    assert res[0] == res[1]
