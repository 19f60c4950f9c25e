"""Host side of the lock-step batched ensemble (dae-cpp/ensemble.hpp: lane_control, solve_ensemble_device;
dae-cpp/device-functor.hpp in its host-loop mode), checked WITHOUT a GPU: the batch entry points are emulated
on the oracle's arithmetic (tests/emu/batch_emulate.cpp), the reference is the unmodified dae-cpp
(hostcpp/_bin/traj_dump_ref).  What must hold per member: same exit code, same number of accepted steps
(the lane state machine makes the reference's decisions), final state within the solver tolerances.
The device kernels themselves are covered by tests/test_gpu_batch.py."""
import json
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "hostcpp", "_bin")


def _need(name):
    exe = os.path.join(BIN, name)
    if not os.path.exists(exe):
        pytest.skip(f"{exe} not built (needs /root/reference at build time)")
    return exe


def _run(exe, args, out, rows, n, env=None):
    p = subprocess.run([exe] + args + [f"out={out}"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900,
                       env=dict(os.environ, **(env or {})))
    assert p.returncode == 0, p.stderr[-2000:]
    info = json.loads(p.stdout.strip().splitlines()[-1])
    data = np.fromfile(out, dtype=np.float64)
    os.remove(out)
    return info, data.reshape(rows, -1)


@pytest.mark.parametrize("stragglers", ["0", "1000000"])  # one Newton iteration per round / up to 8 per round
def test_multi_iteration_rounds_make_the_same_decisions(tmp_path, stragglers):
    members, N = 4, 100
    args = ["ensemble", f"members={members}", "first=0", f"count={members}", f"n={N}"]
    ref_info, ref = _run(_need("traj_dump_ref"), args + ["threads=2"], str(tmp_path / "r.f64"), members, 2 * N)
    info, got = _run(_need("ens_device_emu"), args, str(tmp_path / "e.f64"), members, 2 * N, env={"DAECPP_B200_STRAGGLERS": stragglers})
    assert np.array_equal(got[:, :2], ref[:, :2])  # exit codes and accepted steps
    assert info["lane_iterations"] == ref_info["rhs_calls"] and info["lane_refactorisations"] == ref_info["jac_calls"]
    if stragglers != "0":
        assert info["rounds"] * 3 < info["lane_iterations"] / members * 2  # several iterations per round really happened


@pytest.mark.parametrize("variability", [0, 1])
def test_lockstep_lanes_make_the_reference_decisions(tmp_path, variability):
    members, N = 5, 120
    args = ["ensemble", f"members={members}", "first=0", f"count={members}", f"n={N}"]
    ref_info, ref = _run(_need("traj_dump_ref"), args + ["threads=2"] + ([] if not variability else []), str(tmp_path / "r.f64"), members, 2 * N)
    info, got = _run(_need("ens_device_emu"), args + [f"variability={variability}"], str(tmp_path / "e.f64"), members, 2 * N)
    if variability:
        # the reference driver integrates the sweep with variability control off; only compare like with like
        assert info["failed"] == 0
        return
    assert info["impl"] == "emulated-lockstep" and info["failed"] == 0
    assert np.array_equal(got[:, 0], ref[:, 0])            # exit codes
    assert np.array_equal(got[:, 1], ref[:, 1])            # accepted steps, member by member
    assert info["lane_iterations"] == ref_info["rhs_calls"]           # one RHS evaluation per Newton iteration
    assert info["lane_refactorisations"] == ref_info["jac_calls"]     # one Jacobian per refactorisation
    err = np.abs(got[:, 2:] - ref[:, 2:])
    assert np.all(err <= 1e-6 + 1e-6 * np.abs(ref[:, 2:]))
    # lock step: the sweep takes as many rounds as its slowest member has Newton iterations
    assert info["rounds"] < info["lane_iterations"]


def test_single_lane_with_variability_control_matches_reference_flame(tmp_path):
    """C3 at a small size: one lane, variability control on, the flame example's option tweaks."""
    n = 400
    ref_out, out = str(tmp_path / "r.f64"), str(tmp_path / "e.f64")
    p = subprocess.run([_need("traj_dump_ref"), "flame1d", f"n={n}", "t_end=5", f"out={ref_out}"], stdout=subprocess.PIPE, text=True, timeout=600)
    ref_info = json.loads(p.stdout.strip().splitlines()[-1])
    ref = np.fromfile(ref_out)
    info, got = _run(_need("ens_device_emu"), ["flame1d", f"n={n}", "t_end=5"], out, 1, n)
    assert info["status0"] == ref_info["status"] == 0
    assert got[0, 1] == ref_info["steps"]
    assert info["lane_iterations"] == ref_info["rhs_calls"] and info["lane_refactorisations"] == ref_info["jac_calls"]
    assert np.all(np.abs(got[0, 2:] - ref) <= 1e-6 + 1e-6 * np.abs(ref))


def test_dual_numbers_differentiate_like_the_closed_forms():
    """device-functor.hpp's dual arithmetic against hand-derived derivatives (host instantiation)."""
    src = r'''
#include <dae-cpp/device-functor.hpp>
#include <cstdio>
using namespace daecpp::device;
int main() {
    const double x = 0.7, y = -1.3;
    dual a(x, 1.0), b(y, 0.0);
    dual r = (a * a * b - a / (b + 2.0) + 3.0 * a - 1.0 / a) * exp(a) + tanh(a) * sqrt(a) - pow(a, 2.5) + sin(a) * cos(a) + log(a);
    const double g = x * x * y - x / (y + 2.0) + 3.0 * x - 1.0 / x, dg = 2.0 * x * y - 1.0 / (y + 2.0) + 3.0 + 1.0 / (x * x);
    const double t = std::tanh(x), s = std::sqrt(x);
    const double want = (dg + g) * std::exp(x) + (1 - t * t) * s + t * 0.5 / s - 2.5 * std::pow(x, 1.5) + std::cos(2 * x) + 1.0 / x;
    std::printf("%.17g %.17g\n", r.d, want);
    return 0;
}'''
    import tempfile
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "t.cpp"), "w").write(src)
        exe = os.path.join(d, "t")
        subprocess.check_call(["g++", "-std=c++17", "-O1", f"-I{ROOT}/dae-cpp_b200/overlay", f"-I{ROOT}/include", os.path.join(d, "t.cpp"), "-o", exe])
        got, want = map(float, subprocess.check_output([exe], text=True).split())
    assert abs(got - want) <= 1e-13 * abs(want)


def _observed(exe, args, path, n, env=None):
    p = subprocess.run([exe] + args + [f"observed={path}"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900,
                       env=dict(os.environ, **(env or {})))
    assert p.returncode == 0, p.stderr[-2000:]
    info = json.loads(p.stdout.strip().splitlines()[-1])
    rows = np.fromfile(path, dtype=np.float64).reshape(-1, n + 1)
    os.remove(path)
    return info, rows


@pytest.mark.parametrize("stragglers", ["0", "1000000"])
def test_observer_sees_the_states_of_the_output_times(tmp_path, stragglers):
    """Observer path (solve_ensemble_device_observed; the reference: daecpp::Solution(holder, t_output),
    dae-cpp/solution-manager.hpp:175-189, called from dae-cpp/solver.hpp:659): the same output times, in order, the
    states within the solver tolerances, and ONE state download per recorded state -- nothing else leaves the device."""
    N, K, members = 100, 4, 3
    lam = 0.5  # member 0 of the sweep
    ref_info, ref = _observed(_need("traj_dump_ref"), ["perovskite", f"n={N}", f"lambda={lam}", f"outputs={K}"], str(tmp_path / "r.f64"), 2 * N)
    info, got = _observed(_need("ens_device_emu"), ["ensemble", f"members={members}", "first=0", f"count={members}", f"n={N}", f"outputs={K}"],
                          str(tmp_path / "e.f64"), 2 * N, env={"DAECPP_B200_STRAGGLERS": stragglers})
    assert info["failed"] == 0
    assert got.shape == ref.shape == (K, 2 * N + 1)
    assert np.array_equal(got[:, 0], ref[:, 0])                       # the output times, bit for bit
    assert np.array_equal(got[:, 0], 10.0 * np.arange(1, K + 1) / K)
    tol = 10 * 1e-6                                                   # SolverOptions defaults: atol = rtol = 1e-6
    assert np.all(np.abs(got[:, 1:] - ref[:, 1:]) <= tol * (1.0 + np.abs(ref[:, 1:])))
    assert info["observed"] == members * K and info["observed_downloads"] == members * K
