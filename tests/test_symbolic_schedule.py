"""CPU checks of the once-per-pattern symbolic analysis (dae-cpp_b200/csrc/symbolic.cpp): the
schedule it emits is walked by tests/emu/mf_emulate.cpp with the same per-front code the CUDA
kernels use, and the solution is held to the linear-solve bar of SURVEY.md 8(c)."""
import numpy as np
import pytest

import emu_lib
import oracle_lib as O
import problems as P

EMPTY = (np.zeros(0), np.zeros(0, np.int32), np.zeros(0, np.int32))


@pytest.fixture(scope="module")
def emu():
    return emu_lib.load()


def _check(emu, n, J, M, alpha, budget, leaf=0, b=None):
    csc = O.assemble_jb(O.port, n, J, M, alpha)
    b = np.full(n, 1.24) if b is None else b
    rc, x, st = emu_lib.solve(emu, n, csc, b, leaf, budget)
    assert rc == 0, (rc, st)
    eta, _ = O.backward_error(n, csc, x, b)
    assert eta <= 1e-12, (eta, st)
    return st


BUDGETS = [0, 600, 2048, 16 * 1024, 96 * 1024]


@pytest.mark.parametrize("budget", BUDGETS)
@pytest.mark.parametrize("n", [1, 2, 3, 17, 1000, 30000])
def test_flame1d_all_budgets(emu, n, budget):
    u = P.flame1d_state(n) if n > 1 else np.array([0.5])
    J = P.flame1d_jacobian(u) if n > 1 else (np.array([-3.0]), np.zeros(1, np.int32), np.zeros(1, np.int32))
    st = _check(emu, n, J, P.identity_mass(n), 50.0, budget)
    if budget == 0:
        assert st["n_units"] == 0 and st["n_big"] == st["n_fronts"]
    if budget >= 2048:
        assert st["n_big"] == 0


@pytest.mark.parametrize("budget", BUDGETS)
@pytest.mark.parametrize("N", [50, 1000])
def test_perovskite_all_budgets(emu, N, budget):
    x = P.perovskite_state(N)
    _check(emu, 2 * N, P.perovskite_jacobian(x, N), P.perovskite_mass(N), 100.0, budget)


@pytest.mark.parametrize("budget", [0, 4096, 96 * 1024])
@pytest.mark.parametrize("m,leaf", [(4, 0), (33, 0), (64, 2), (96, 0)])
def test_grid2d_units_and_big_fronts(emu, m, leaf, budget):
    st = _check(emu, m * m, P.grid2d_probe_matrix(m), EMPTY, 0.0, budget, leaf)
    if budget == 4096 and m >= 33:
        assert st["n_big"] > 0 and st["n_units"] > 0 and st["n_groups"] >= 2  # mixed schedule


@pytest.mark.parametrize("m,budget", [(96, 4096), (128, 2048), (160, 8192), (200, -1)])
def test_units_beside_big_fronts_of_their_stage_are_independent(emu, m, budget):
    """The CUDA path runs a launch of units beside the level of big fronts behind it when the schedule flags the pair
    (LaunchGroup::same_stage).  The scalar walk takes every flagged pair in the opposite order, in the factorisation and
    in both sweeps: the answer must not care.  (n_groups carries the number of swapped pairs in its fraction.)"""
    st = _check(emu, m * m, P.grid2d_probe_matrix(m), EMPTY, 0.0, budget)
    swapped = round((st["n_groups"] - int(st["n_groups"])) * 1000)
    if budget >= 0:
        assert st["n_big"] > 0 and swapped >= 1, st


@pytest.mark.parametrize("budget", [0, 1024, 96 * 1024])
def test_disconnected_and_dense_blocks(emu, budget):
    r = np.random.default_rng(3)
    n = 100
    A = np.zeros((n, n))
    A[:40, :40] = r.standard_normal((40, 40)) + 40 * np.eye(40)
    A[40:80, 40:80] = r.standard_normal((40, 40)) + 40 * np.eye(40)
    A[80:, 80:] = np.diag(r.uniform(1, 2, 20))
    ii, jj = np.nonzero(A)
    _check(emu, n, (A[ii, jj], ii.astype(np.int32), jj.astype(np.int32)), EMPTY, 0.0, budget, b=r.standard_normal(n))


def test_identity_packs_roots_into_few_units(emu):
    n = 50000
    st = _check(emu, n, P.identity_mass(n), EMPTY, 0.0, 96 * 1024)
    assert st["n_units"] < 200  # parentless roots are binned, not one unit each


def test_unsymmetric_pattern(emu):
    # lower-bidiagonal + a few far upper entries: pattern(A+A^T) drives the analysis
    n = 500
    k = np.arange(n, dtype=np.int32)
    A = np.concatenate([np.full(n, 4.0), np.full(n - 1, -1.0), np.full(5, 0.5)])
    i = np.concatenate([k, k[1:], np.array([0, 10, 20, 30, 40], np.int32)])
    j = np.concatenate([k, k[:-1], np.array([400, 410, 420, 430, 440], np.int32)])
    for budget in (0, 2048, 96 * 1024):
        _check(emu, n, (A, i, j), EMPTY, 0.0, budget)
