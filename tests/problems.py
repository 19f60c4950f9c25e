"""Synthetic inputs for the hot path: the triplet streams the user functors of the BASELINE
configs hand to the solver.  numpy only; deterministic.

C2  perovskite  -- same entries, same emission order (row by row) as
                   /root/reference/examples/perovskite_model/perovskite.cpp:132-186
C3  flame-1D    -- u_t = D u_xx + u^2 - u^3, 3-point stencil, zero-flux ends
C4  grid-2D     -- u_t = Lap u + u^2 - u^3 on an m x m grid, 5-point stencil, Neumann edges
"""
import numpy as np


def perovskite_jacobian(x, N, L=1.0, lam=1.0):
    h = L / (N - 1)
    invh2 = 1.0 / (h * h)
    invlam2 = 1.0 / (lam * lam)
    A, I, J = [], [], []

    def e(i, j, v):
        I.append(i); J.append(j); A.append(v)

    for i in range(2 * N):
        if i == 0:
            e(i, 0, (-1.0 + 0.5 * (x[N + 1] - x[N])) * invh2)
            e(i, 1, (1.0 + 0.5 * (x[N + 1] - x[N])) * invh2)
            e(i, N, -0.5 * (x[0] + x[1]) * invh2)
            e(i, N + 1, 0.5 * (x[0] + x[1]) * invh2)
        elif i < N - 1:
            e(i, i - 1, (1.0 - 0.5 * (x[N + i] - x[N + i - 1])) * invh2)
            e(i, i, (-2.0 + 0.5 * (x[N + i + 1] - 2.0 * x[N + i] + x[N + i - 1])) * invh2)
            e(i, i + 1, (1.0 + 0.5 * (x[N + i + 1] - x[N + i])) * invh2)
            e(i, N + i - 1, 0.5 * (x[i] + x[i - 1]) * invh2)
            e(i, N + i, -0.5 * (x[i + 1] + 2.0 * x[i] + x[i - 1]) * invh2)
            e(i, N + i + 1, 0.5 * (x[i + 1] + x[i]) * invh2)
        elif i == N - 1:
            e(i, i - 1, (1.0 - 0.5 * (x[2 * N - 1] - x[2 * N - 2])) * invh2)
            e(i, i, (-1.0 - 0.5 * (x[2 * N - 1] - x[2 * N - 2])) * invh2)
            e(i, N + i - 1, 0.5 * (x[N - 1] + x[N - 2]) * invh2)
            e(i, N + i, -0.5 * (x[N - 1] + x[N - 2]) * invh2)
        elif i == N:
            e(i, N, 1.0)
        elif i < 2 * N - 1:
            e(i, i - N, invlam2)
            e(i, i - 1, invh2)
            e(i, i, -2.0 * invh2)
            e(i, i + 1, invh2)
        else:
            e(i, 2 * N - 1, 1.0)
    return np.array(A, np.float64), np.array(I, np.int32), np.array(J, np.int32)


def perovskite_mass(N):
    k = np.arange(N, dtype=np.int32)
    return np.ones(N), k.copy(), k.copy()


def perovskite_state(N, seed=0):
    """A smooth, physically plausible state: P ~ 1 +- 0.3, Phi a ramp with wiggles."""
    r = np.random.default_rng(seed)
    s = np.linspace(0.0, 1.0, N)
    P = 1.0 + 0.3 * np.sin(3.0 * s + r.uniform()) * r.uniform(0.5, 1.0)
    Phi = -1.0 + 2.0 * s + 0.1 * np.sin(7.0 * s)
    return np.concatenate([P, Phi])


def identity_mass(n):
    k = np.arange(n, dtype=np.int32)
    return np.ones(n), k.copy(), k.copy()


def flame1d_state(n, delta=1e-4):
    s = np.linspace(0.0, 1.0, n)
    return delta + (1.0 - delta) * 0.5 * (1.0 - np.tanh((s - 0.1) / 0.02))


def flame1d_jacobian(u, D=1e-3):
    """Row-by-row (i,i-1),(i,i),(i,i+1) as JacobianMatrixShape::add_element(i,{i-1,i,i+1}) emits."""
    n = len(u)
    h = 1.0 / (n - 1)
    c = D / (h * h)
    react = 2.0 * u - 3.0 * u * u
    A = np.empty((n, 3)); I = np.empty((n, 3), np.int32); J = np.empty((n, 3), np.int32)
    k = np.arange(n, dtype=np.int32)
    I[:] = k[:, None]
    J[:, 0] = k - 1; J[:, 1] = k; J[:, 2] = k + 1
    A[:, 0] = c; A[:, 2] = c
    A[:, 1] = -2.0 * c + react
    A[0, 1] = -c + react[0]; A[-1, 1] = -c + react[-1]
    keep = np.ones((n, 3), bool); keep[0, 0] = False; keep[-1, 2] = False
    return A[keep].copy(), I[keep].copy(), J[keep].copy()


def flame1d_rhs(u, D=1e-3):
    n = len(u)
    h = 1.0 / (n - 1)
    c = D / (h * h)
    f = np.empty(n)
    f[1:-1] = c * (u[2:] - 2 * u[1:-1] + u[:-2])
    f[0] = c * (u[1] - u[0]); f[-1] = c * (u[-2] - u[-1])
    return f + u * u - u ** 3


def grid2d_state(m, delta=1e-4):
    r, c = np.meshgrid(np.arange(m), np.arange(m), indexing="ij")
    d2 = ((r - m / 2) ** 2 + (c - m / 2) ** 2) / float(m * m)
    return np.where(d2 < 0.01, 1.0, delta).ravel()


def grid2d_jacobian(u, m):
    """Per row k=r*m+c: (k,k-m),(k,k-1),(k,k),(k,k+1),(k,k+m) where they exist."""
    n = m * m
    h = 1.0 / (m - 1)
    cc = 1.0 / (h * h)
    k = np.arange(n, dtype=np.int32)
    r = k // m; c = k % m
    A = np.empty((n, 5)); I = np.repeat(k[:, None], 5, 1); J = np.empty((n, 5), np.int32)
    off = np.array([-m, -1, 0, 1, m], np.int32)
    J[:] = k[:, None] + off[None, :]
    keep = np.ones((n, 5), bool)
    keep[:, 0] = r > 0; keep[:, 1] = c > 0; keep[:, 3] = c < m - 1; keep[:, 4] = r < m - 1
    nb = keep[:, 0].astype(float) + keep[:, 1] + keep[:, 3] + keep[:, 4]
    A[:] = cc
    A[:, 2] = -nb * cc + 2.0 * u - 3.0 * u * u
    return A[keep].copy(), I[keep].astype(np.int32).copy(), J[keep].copy()


def grid2d_probe_matrix(m):
    """The pure linear-solve microbench matrix of SURVEY 8(d): diag -4-10-0.1*((131r+17c) mod 10), off-diag 1."""
    n = m * m
    A, I, J = grid2d_jacobian(np.zeros(n), m)
    k = np.arange(n)
    r = k // m; c = k % m
    d = -4.0 - 10.0 - 0.1 * ((131 * r + 17 * c) % 10)
    diag = I == J
    A[:] = 1.0
    A[diag] = d[I[diag]]
    return A, I, J


def random_triplets(n, nT, seed, dup_frac=0.3):
    r = np.random.default_rng(seed)
    i = r.integers(0, n, nT).astype(np.int32)
    j = r.integers(0, n, nT).astype(np.int32)
    nd = int(nT * dup_frac)
    if nd and nT > 1:
        src = r.integers(0, nT, nd)
        dst = r.integers(0, nT, nd)
        i[dst] = i[src]; j[dst] = j[src]
    A = r.standard_normal(nT) * 10.0 ** r.integers(-8, 8, nT)
    return A, i, j
