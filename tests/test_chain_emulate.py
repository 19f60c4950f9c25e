"""CPU walk over the block-chain plan (tests/emu/chain_emulate.cpp): detection, ordering, slot map
and the three-level-per-CTA elimination hierarchy of dae-cpp_b200/csrc/chain_ops.hpp, checked by the
normwise backward error of the solution (the bar of SURVEY.md 8(c)2: eta <= 1e-12)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
import scipy.sparse as sp

import problems as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(ROOT, "tests", "emu", "libchainemu.so")
_SRC = [os.path.join(ROOT, "tests", "emu", "chain_emulate.cpp"), os.path.join(ROOT, "dae-cpp_b200", "csrc", "chain_host.cpp")]
_HDR = [os.path.join(ROOT, "dae-cpp_b200", "csrc", h) for h in ("chain.hpp", "chain_ops.hpp")]


@pytest.fixture(scope="module")
def lib():
    newest = max(os.path.getmtime(p) for p in _SRC + _HDR)
    if not os.path.exists(_SO) or os.path.getmtime(_SO) < newest:
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", *_SRC, "-o", _SO])
    L = C.CDLL(_SO)
    f64 = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
    i32 = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
    L.chainemu_solve.argtypes = [C.c_int32, i32, i32, f64, f64, f64, C.c_int32, f64]
    return L


def run(lib, A, b, max_b=3):
    A = sp.csc_matrix(A)
    A.sort_indices()
    n = A.shape[0]
    x = np.zeros(n)
    st = np.zeros(8)
    rc = lib.chainemu_solve(n, A.indptr.astype(np.int32), A.indices.astype(np.int32), np.ascontiguousarray(A.data),
                            np.ascontiguousarray(b), x, max_b, st)
    if rc:
        return rc, None, st
    r = b - A @ x
    eta = np.abs(r).max() / (abs(A).sum(axis=1).max() * np.abs(x).max() + np.abs(b).max())
    return rc, eta, st


@pytest.mark.parametrize("n", [1, 2, 5, 15, 16, 17, 255, 2047, 2048, 2049, 5000, 100000])
def test_tridiagonal_sizes(lib, n):
    rng = np.random.default_rng(n)
    d = -(2.5 + rng.random(n))
    if n > 1:
        A = sp.diags([1 + 0.1 * rng.random(n - 1), d, 1 + 0.1 * rng.random(n - 1)], [-1, 0, 1], format="csc")
    else:
        A = sp.csc_matrix(np.array([[d[0]]]))
    rc, eta, st = run(lib, A, rng.standard_normal(n))
    assert rc == 0 and st[0] == 1 and st[3] == 1
    assert st[2] == (1 if n <= 2048 else 2)
    assert eta <= 1e-14


def test_three_grid_levels(lib):
    n = 2048 * 2048 + 5
    rng = np.random.default_rng(1)
    A = sp.diags([np.ones(n - 1), -(2.5 + rng.random(n)), np.ones(n - 1)], [-1, 0, 1], format="csc")
    rc, eta, st = run(lib, A, rng.standard_normal(n))
    assert rc == 0 and st[2] == 3 and eta <= 1e-14


def test_flame1d_c3_shape(lib):
    n = 200000
    u = P.flame1d_state(n)
    Av, Ai, Aj = P.flame1d_jacobian(u)
    A = sp.csc_matrix((Av, (Ai, Aj)), shape=(n, n)) - 50.0 * sp.eye(n)
    rc, eta, st = run(lib, A, np.full(n, 1.24))
    assert rc == 0 and st[0] == 1 and eta <= 1e-12


@pytest.mark.parametrize("N", [5, 50, 300, 4000])
def test_perovskite_is_a_chain_of_2x2_blocks(lib, N):
    n = 2 * N
    x = P.perovskite_state(N)
    Av, Ai, Aj = P.perovskite_jacobian(x, N)
    Mv, Mi, Mj = P.perovskite_mass(N)
    A = sp.csc_matrix((Av, (Ai, Aj)), shape=(n, n)) - 100.0 * sp.csc_matrix((Mv, (Mi, Mj)), shape=(n, n))
    rc, eta, st = run(lib, A, np.full(n, 1.24))
    assert rc == 0 and st[0] == 2 and st[3] == 0  # interleaving permutation found
    assert eta <= 1e-14


@pytest.mark.parametrize("n", [7, 100, 3001])
def test_block3_band_with_padding(lib, n):
    rng = np.random.default_rng(n)
    A = sp.lil_matrix((n, n))
    for i in range(n):
        for j in range(max(0, i - 4), min(n, i + 5)):
            if rng.random() < 0.6 and abs(i // 3 - j // 3) <= 1:
                A[i, j] = rng.standard_normal()
        A[i, i] = 8 + rng.random()
    rc, eta, st = run(lib, A, rng.standard_normal(n))
    assert rc == 0 and st[0] == 3 and eta <= 1e-14


def test_pivoting_inside_a_block(lib):
    # zero diagonal inside every 2x2 block: only the in-block row exchange makes this factorisable
    m = 700
    n = 2 * m
    rng = np.random.default_rng(3)
    A = sp.lil_matrix((n, n))
    for k in range(m):
        A[2 * k, 2 * k + 1] = 3 + rng.random()
        A[2 * k + 1, 2 * k] = -3 - rng.random()
        A[2 * k + 1, 2 * k + 1] = 0.5
        if k + 1 < m:
            A[2 * k, 2 * k + 2] = 0.3
            A[2 * k + 3, 2 * k + 1] = -0.2
    rc, eta, st = run(lib, A, rng.standard_normal(n))
    assert rc == 0 and st[0] == 2 and eta <= 1e-14


def test_permuted_tridiagonal_is_recovered(lib):
    n = 3000
    rng = np.random.default_rng(0)
    p = rng.permutation(n)
    A = sp.csc_matrix(sp.diags([np.ones(n - 1), -3 * np.ones(n), np.ones(n - 1)], [-1, 0, 1]))[p][:, p]
    rc, eta, st = run(lib, A, rng.standard_normal(n))
    assert rc == 0 and st[0] == 1 and st[3] == 0 and eta <= 1e-14


def test_disconnected_chains(lib):
    blocks = [sp.diags([np.ones(k - 1), -3 * np.ones(k), np.ones(k - 1)], [-1, 0, 1]) for k in (40, 7, 900)]
    A = sp.block_diag(blocks, format="csc")
    rc, eta, st = run(lib, A, np.ones(A.shape[0]))
    assert rc == 0 and eta <= 1e-14


def test_not_a_chain_is_declined(lib):
    A = sp.random(200, 200, 0.1, random_state=1) + sp.eye(200)
    assert run(lib, A, np.ones(200))[0] == 1
    m = 12  # 2-D grid: bandwidth m
    assert run(lib, sp.kronsum(sp.diags([1, -2, 1], [-1, 0, 1], shape=(m, m)), sp.diags([1, -2, 1], [-1, 0, 1], shape=(m, m))),
               np.ones(m * m))[0] == 1


def test_missing_diagonal_goes_to_general_path(lib):
    n = 50
    A = sp.diags([np.ones(n - 1), np.ones(n - 1)], [-1, 1], format="csc")  # no diagonal at all, block order 1
    assert run(lib, A, np.ones(n), max_b=1)[0] == 1
