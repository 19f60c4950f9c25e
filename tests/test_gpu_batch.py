"""Lock-step batched ensemble on the device (include/daecpp_b200.h: daecpp_b200_batch_*).

1. A lane of a batch is BIT-IDENTICAL to the same Newton iteration run through a single-system context on the
   same inputs (same kernels, lane = blockIdx.y): iterate, dx, flags -- with refactorisation, with reused
   factors, after an accepted step (history ring), after a rollback, with idle lanes in between.
2. Whole ensembles through dae-cpp/ensemble.hpp: solve_ensemble_device (device functors, per-lane adaptive
   stepping) against the unmodified reference member by member: exit code, accepted steps +-2 %, final state
   within 10 x (atol, rtol)  (BASELINE.json north_star; SURVEY.md 8(c) definition 3).
3. The device functors (dae-cpp/device-functor.hpp) against the host triplet streams of tests/problems.py.
"""
import json
import os
import subprocess

import numpy as np
import pytest

import problems as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "hostcpp", "_bin")
pytestmark = pytest.mark.gpu


def _need(name):
    exe = os.path.join(BIN, name)
    if not os.path.exists(exe):
        pytest.skip(f"{exe} not built (needs /root/reference at build time)")
    return exe


def _perovskite_case(N, lanes, seed):
    r = np.random.default_rng(seed)
    xs = [P.perovskite_state(N, seed + l) for l in range(lanes)]
    Js = [P.perovskite_jacobian(x, N, lam=0.5 + 0.3 * l) for l, x in enumerate(xs)]
    M = P.perovskite_mass(N)
    fs = [r.standard_normal(2 * N) for _ in range(lanes)]
    return 2 * N, xs, Js, M, fs


def _flame_case(n, lanes, seed):
    r = np.random.default_rng(seed)
    xs = [np.clip(P.flame1d_state(n) + 0.01 * r.standard_normal(n), 1e-6, 1.0) for _ in range(lanes)]
    Js = [P.flame1d_jacobian(x) for x in xs]
    M = P.identity_mass(n)
    fs = [P.flame1d_rhs(x) for x in xs]
    return n, xs, Js, M, fs


@pytest.mark.parametrize("case", ["perovskite_1tile", "perovskite_3tiles", "flame_3tiles"])
def test_batch_lane_bit_identical_to_single_context(case):
    from gpu_lib import binding as B
    lanes = 4
    if case == "perovskite_1tile":
        n, xs, Js, M, fs = _perovskite_case(200, lanes, 1)
    elif case == "perovskite_3tiles":
        n, xs, Js, M, fs = _perovskite_case(700, lanes, 2)
    else:
        n, xs, Js, M, fs = _flame_case(5000, lanes, 3)
    atol = rtol = 1e-6
    lib = B.load()
    b = B.Batch(n, lanes, 0, lib)
    b.set_pattern(Js[0][1], Js[0][2], M[1], M[2])
    b.set_mass_values(M[0])
    b.set_state(np.stack(xs))
    singles = []
    for l in range(lanes):
        c = B.Context(n, 0, lib)
        c.set_state(xs[l]); c.set_mass(M); c.set_jacobian(Js[l])
        singles.append(c)

    def single_iter(l, f, refactor, order, dt):
        c = singles[l]
        c.residual(f, order, dt)  # uploads f (and forms b; the resident iteration forms it again)
        rc, flags = c.iteration_resident(refactor, order, dt, atol, rtol, allow=(B.EDECOMP,))
        return rc, flags

    def compare(active):
        xk, dx = b.get_vector("xk"), b.get_vector("dx")
        for l in active:
            assert xk[l].tobytes() == singles[l].get_vector("xk").tobytes(), f"lane {l}: iterate differs"
            assert dx[l].tobytes() == singles[l].get_vector("dx").tobytes(), f"lane {l}: dx differs"

    # round 1: every lane refactors, BDF-1, lane-dependent step size
    dts = [np.array([1e-2 * (l + 1), 0, 0, 0.0]) for l in range(lanes)]
    b.set_jacobian_values(np.stack([J[0] for J in Js])); b.set_rhs_values(np.stack(fs))
    rc, res = b.round([dict(mode=B.LANE_REFACTOR, order=1, dt=dts[l]) for l in range(lanes)], atol, rtol)
    for l in range(lanes):
        rc1, fl1 = single_iter(l, fs[l], 1, 1, dts[l])
        assert res[l][1] == rc1 == 0 and res[l][0] == fl1
    compare(range(lanes))
    assert b.stat("chain_block") == singles[0].stat("chain_block") > 0
    # round 2: lanes 0 and 2 reuse their factors, lane 1 refactors with new values, lane 3 sits out
    fs2 = [f * 0.5 for f in fs]
    J1 = Js[1][0] * 1.01
    b.set_rhs_values(np.stack(fs2)); b.set_jacobian_values(J1, 1, 1)
    cmds = [dict(mode=B.LANE_SOLVE, order=1, dt=dts[0]), dict(mode=B.LANE_REFACTOR, order=1, dt=dts[1]),
            dict(mode=B.LANE_SOLVE, order=1, dt=dts[2]), dict(mode=B.LANE_IDLE)]
    x3_before = b.get_vector("xk")[3].copy()
    rc, res = b.round(cmds, atol, rtol)
    singles[1].set_jacobian((J1, Js[1][1], Js[1][2]))
    for l, refac in ((0, 0), (1, 1), (2, 0)):
        rc1, fl1 = single_iter(l, fs2[l], refac, 1, dts[l])
        assert res[l][1] == rc1 == 0 and res[l][0] == fl1
    compare((0, 1, 2))
    assert b.get_vector("xk")[3].tobytes() == x3_before.tobytes() and res[3] == (0, 0, 0.0)
    # round 3: lane 0 accepts its step and goes on at BDF-2 with the variability measured in the update,
    # lane 2 rolls back, lanes 1 and 3 idle
    dt2 = np.array([2e-2, dts[0][0], 0, 0.0])
    b.set_rhs_values(fs[0], 0, 1); b.set_rhs_values(fs[2], 2, 1)
    b.set_jacobian_values(Js[0][0], 0, 1); b.set_jacobian_values(Js[2][0], 2, 1)
    cmds = [dict(mode=B.LANE_REFACTOR, pre=B.LANE_ACCEPT, order=2, dt=dt2), dict(mode=B.LANE_IDLE),
            dict(mode=B.LANE_REFACTOR, pre=B.LANE_ROLLBACK, order=1, dt=dts[2]), dict(mode=B.LANE_IDLE)]
    rc, res = b.round(cmds, atol, rtol, variability_on=True, var_tol=1e-4)
    singles[0].accept_step(False); singles[2].rollback()
    x0_old = singles[0].get_vector("x0")
    rc1, fl1 = single_iter(0, fs[0], 1, 2, dt2)
    assert res[0][1] == rc1 == 0 and res[0][0] == fl1
    rc1, fl1 = single_iter(2, fs[2], 1, 1, dts[2])
    assert res[2][1] == rc1 == 0 and res[2][0] == fl1
    compare((0, 2))
    for k in range(4):
        assert b.get_vector(f"x{k}", 0, 1)[0].tobytes() == singles[0].get_vector(f"x{k}").tobytes()
    # the variability the update reported is what accept_step measures afterwards (solver.hpp:703-717)
    assert res[0][2] == singles[0].accept_step(True, 1e-4)
    assert np.array_equal(singles[0].get_vector("x1"), x0_old)
    for c in singles:
        c.close()
    b.close()


def test_multi_iteration_round_equals_single_iteration_rounds():
    """One round of up to 4 Newton iterations per lane (include/daecpp_b200.h: daecpp_b200_lane_cmd.max_iterations) against 4
    rounds of one: same iterates bit for bit, lanes stop on the device where the reference's loop stops (converged)."""
    from gpu_lib import binding as B
    import ctypes as C
    lib = B.load()
    exe_lib = os.path.join(BIN, "libensdevice.so")
    if not os.path.exists(exe_lib):
        pytest.skip("hostcpp/_bin/libensdevice.so not built")
    # device functors come from the perovskite driver library: drive two whole sweeps, one regime each
    ens = C.CDLL(exe_lib)
    ens.ensdev_perovskite.argtypes = [C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_double, C.c_int32, C.c_void_p, C.c_void_p]
    N, members = 300, 6
    outs, stats = [], []
    for stragglers in ("0", "1000000"):
        os.environ["DAECPP_B200_STRAGGLERS"] = stragglers
        out = np.zeros((members, 2 * N + 2)); st = np.zeros(16)
        ens.ensdev_perovskite(members, 0, members, 1, N, 10.0, 0, out.ctypes.data, st.ctypes.data)
        outs.append(out); stats.append(st)
    os.environ.pop("DAECPP_B200_STRAGGLERS")
    assert outs[0].tobytes() == outs[1].tobytes()            # status, steps and final states identical
    assert stats[0][1] == stats[1][1] and stats[0][2] == stats[1][2]  # lane iterations / refactorisations
    assert stats[1][0] * 2 < stats[0][0]                     # in far fewer rounds


def test_batch_refuses_what_it_cannot_take():
    from gpu_lib import binding as B
    lib = B.load()
    m = 24
    J = P.grid2d_jacobian(P.grid2d_state(m), m)
    M = P.identity_mass(m * m)
    b = B.Batch(m * m, 2, 0, lib)
    assert b.set_pattern(J[1], J[2], M[1], M[2], allow=(B.EINVAL,)) == B.EINVAL  # a 2-D grid is no block chain
    b.close()
    n = 600
    x = P.flame1d_state(n)
    A, i, j = P.flame1d_jacobian(x)
    b = B.Batch(n, 2, 0, lib)
    assert b.set_pattern(np.concatenate([i, i[:5]]), np.concatenate([j, j[:5]]), *P.identity_mass(n)[1:], allow=(B.EINVAL,)) == B.EINVAL  # duplicates
    b.close()


def test_singular_lane_reports_decomposition_failure_and_leaves_the_others_alone():
    from gpu_lib import binding as B
    lib = B.load()
    n, lanes = 3000, 3
    x = P.flame1d_state(n)
    A, i, j = P.flame1d_jacobian(x)
    M = P.identity_mass(n)
    b = B.Batch(n, lanes, 0, lib)
    b.set_pattern(i, j, M[1], M[2]); b.set_mass_values(M[0]); b.set_state(np.stack([x] * lanes))
    bad = A.copy()
    bad[:] = 0.0  # J = 0 and alpha*M cancelled below: a zero matrix
    Mz = np.zeros(n)
    b.set_mass_values(np.stack([M[0], Mz, M[0]]), 0, 3)
    b.set_jacobian_values(np.stack([A, bad, A])); b.set_rhs_values(np.stack([P.flame1d_rhs(x)] * lanes))
    rc, res = b.round([dict(mode=B.LANE_REFACTOR, order=1, dt=[1e-2, 0, 0, 0])] * lanes, 1e-6, 1e-6)
    assert rc == 0
    assert res[1][1] == B.EDECOMP and res[0][1] == 0 and res[2][1] == 0
    xk = b.get_vector("xk")
    assert xk[0].tobytes() == xk[2].tobytes() and np.all(np.isfinite(xk[0]))
    b.close()


def _members(exe, args, out, members, n):
    p = subprocess.run([exe, "ensemble", f"members={members}", "first=0", f"count={members}", f"n={n}", f"out={out}"] + args,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900)
    assert p.returncode == 0, p.stderr[-3000:]
    info = json.loads(p.stdout.strip().splitlines()[-1])
    data = np.fromfile(out, dtype=np.float64).reshape(members, 2 * n + 2)
    os.remove(out)
    return info, data


@pytest.mark.parametrize("N,members", [(150, 16), (300, 12), (4000, 24)])
def test_lockstep_ensemble_members_match_reference(tmp_path, N, members):
    ref_info, ref = _members(_need("traj_dump_ref"), ["threads=8"], str(tmp_path / "ref.f64"), members, N)
    info, got = _members(_need("ens_device"), [], str(tmp_path / "gpu.f64"), members, N)
    assert info["impl"] == "b200-lockstep" and info["failed"] == 0
    assert np.array_equal(got[:, 0], ref[:, 0])  # exit codes
    assert np.all(np.abs(got[:, 1] - ref[:, 1]) <= np.maximum(0.02 * ref[:, 1], 0.5))  # accepted steps +-2 %
    atol = rtol = 1e-6
    err = np.abs(got[:, 2:] - ref[:, 2:])
    assert np.all(err <= 10.0 * (atol + rtol * np.abs(ref[:, 2:]))), float(err.max())
    assert info["rounds"] < info["lane_iterations"] / 2  # the lanes really shared their rounds


@pytest.mark.parametrize("N,members", [(300, 5), (4000, 6)])
def test_observer_path_downloads_only_the_output_states(tmp_path, N, members):
    """f3: solve_ensemble_device_observed against daecpp::Solution(holder, t_output) of the unmodified
    reference (dae-cpp/solution-manager.hpp:175-189): the same output times bit for bit, the recorded states within
    10x the tolerances, and exactly one lane download per recorded state."""
    K = 5
    def observed(exe, args):
        path = str(tmp_path / "obs.f64")
        p = subprocess.run([exe] + args + [f"observed={path}"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900)
        assert p.returncode == 0, p.stderr[-2000:]
        rows = np.fromfile(path, dtype=np.float64).reshape(-1, 2 * N + 1)
        os.remove(path)
        return json.loads(p.stdout.strip().splitlines()[-1]), rows
    _, ref = observed(_need("traj_dump_ref"), ["perovskite", f"n={N}", "lambda=0.5", f"outputs={K}"])
    info, got = observed(_need("ens_device"), ["ensemble", f"members={members}", "first=0", f"count={members}", f"n={N}", f"outputs={K}"])
    assert info["impl"] == "b200-lockstep" and info["failed"] == 0
    assert got.shape == ref.shape == (K, 2 * N + 1)
    assert np.array_equal(got[:, 0], ref[:, 0])
    err = np.abs(got[:, 1:] - ref[:, 1:])
    assert np.all(err <= 10.0 * (1e-6 + 1e-6 * np.abs(ref[:, 1:]))), float(err.max())
    assert info["observed"] == members * K and info["observed_downloads"] == members * K


@pytest.mark.parametrize("N,members,first", [(300, 10, 0), (4000, 16, 360)])
def test_lockstep_members_bit_identical_to_single_context_solves(tmp_path, N, members, first):
    """Whole trajectories: every lane of the lock-step batch (device functors with the example's closed-form Jacobian,
    -fmad=false) against the SAME member integrated alone by daecpp::solve through the overlay header -- one context,
    host functors: accepted steps, exit code and the final state BIT FOR BIT.  Members 360..375 of the C5 sweep include
    the one whose potential crosses zero at a grid point (the reference's relative test is on a knife edge there)."""
    def run(exe, extra):
        out = str(tmp_path / "o.f64")
        p = subprocess.run([exe, "ensemble", "members=4096", f"first={first}", f"count={members}", f"n={N}", f"out={out}"] + extra,
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900)
        assert p.returncode in (0, 1), p.stderr[-2000:]
        data = np.fromfile(out, dtype=np.float64).reshape(members, 2 * N + 2)
        os.remove(out)
        return data
    single = run(_need("traj_dump"), ["threads=4"])
    batch = run(_need("ens_device"), [])
    assert np.array_equal(batch[:, :2], single[:, :2])
    assert batch[:, 2:].tobytes() == single[:, 2:].tobytes()


def test_single_lane_device_functor_trajectory_matches_reference_golden(tmp_path):
    """C3 (flame-1D, n = 20 000) integrated entirely on the device -- RHS and shape Jacobian by device functors,
    state never on the host -- against the reference's golden trajectory."""
    meta = json.load(open(os.path.join(ROOT, "tests", "golden", "traj_flame1d_20000.json")))
    want = np.fromfile(os.path.join(ROOT, "tests", "golden", "traj_flame1d_20000.f64"))
    out = str(tmp_path / "x.f64")
    p = subprocess.run([_need("ens_device"), "flame1d", "n=20000", "t_end=5", f"out={out}"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    info = json.loads(p.stdout.strip().splitlines()[-1])
    got = np.fromfile(out)
    assert got[0] == meta["status"] == 0
    assert abs(got[1] - meta["steps"]) <= max(0.02 * meta["steps"], 0.5)
    tol = 10.0 * meta["tol"]
    assert np.all(np.abs(got[2:] - want) <= tol + tol * np.abs(want)), float(np.max(np.abs(got[2:] - want)))
    assert info["rounds"] <= info["lane_iterations"] == 401  # (a lone lane runs several Newton iterations per round)


def test_guarded_refinement_single_and_batched():
    """The update fails the relative test only on elements whose |dx| sits below the refinement floor (an element of
    x[0] that is tiny but above the 1e-14 floor of solver.hpp:603): one refinement step runs, in the single-system
    context and in the batch alike, with bit-identical results; a converged / clearly unconverged update takes none."""
    from gpu_lib import binding as B
    import scipy.sparse as sp
    lib = B.load()
    n, lanes = 6000, 3
    u = P.flame1d_state(n)
    u[n // 2] = 3e-13  # passes the 1e-14 floor, so the relative test needs |dx| < 3e-19 there
    A, i, j = P.flame1d_jacobian(u)
    M = P.identity_mass(n)
    dt = np.array([1e-2, 0, 0, 0.0])
    Jb = sp.csr_matrix((A, (i, j)), shape=(n, n)) - sp.identity(n) * (1.0 / dt[0])
    r = np.random.default_rng(5)
    z = 1e-13 * r.standard_normal(n)          # the update the iteration must deliver: everywhere below the floor 1e-11
    f = -(Jb @ z)                              # b = M*dxdt - f = -f  (xk == x[0] at the first iteration of a BDF-1 step)
    c = B.Context(n, 0, lib)
    c.set_state(u); c.set_mass(M); c.set_jacobian((A, i, j))
    c.residual(f, 1, dt)
    rc, flags = c.iteration_resident(1, 1, dt, 1e-6, 1e-6)
    assert rc == 0 and flags == B.NOT_CONVERGED and c.stat("n_refine") == 1
    dx = c.get_vector("dx")
    assert np.max(np.abs(dx - z)) <= 1e-12 * np.max(np.abs(z))
    # the same lane three times in a batch, the middle one with a large update (no refinement for it)
    b = B.Batch(n, lanes, 0, lib)
    b.set_pattern(i, j, M[1], M[2]); b.set_mass_values(M[0]); b.set_state(np.stack([u] * lanes))
    b.set_jacobian_values(np.stack([A] * lanes)); b.set_rhs_values(np.stack([f, 1e6 * f, f]))
    rc, res = b.round([dict(mode=B.LANE_REFACTOR, order=1, dt=dt)] * lanes, 1e-6, 1e-6)
    assert [q[0] for q in res] == [B.NOT_CONVERGED] * 3 and [q[1] for q in res] == [0, 0, 0]
    assert b.stat("refine_rounds") == 1 and b.stat("refined_lanes") == 2
    bdx, bxk = b.get_vector("dx"), b.get_vector("xk")
    assert bdx[0].tobytes() == dx.tobytes() == bdx[2].tobytes()
    assert bxk[0].tobytes() == c.get_vector("xk").tobytes()
    assert np.max(np.abs(bdx[1] - 1e6 * z)) <= 1e-9 * np.max(np.abs(1e6 * z))
    # refinement switched off: the unrefined update is what the plain solve delivers
    c2 = B.Context(n, 0, lib)
    c2.set_option("refine_steps", 0)
    c2.set_state(u); c2.set_mass(M); c2.set_jacobian((A, i, j))
    c2.residual(f, 1, dt)
    rc, flags = c2.iteration_resident(1, 1, dt, 1e-6, 1e-6)
    assert c2.stat("n_refine") == 0 and flags == B.NOT_CONVERGED
    for q in (c, c2):
        q.close()
    b.close()
