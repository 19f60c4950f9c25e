"""Ensemble (C5): members are independent daecpp::solve instances partitioned across
ranks, no data-path collective, one final gather.

CPU (gloo, world_size 2): the partition + gather logic of bench.py gives the same per-member results
whatever the number of ranks (the reference driver integrates the members here -- host logic only).
GPU: every member integrated through the overlay on the device matches the same member integrated by
the unmodified reference: exit code, accepted steps within +-2 %, final state within 10x (atol, rtol).
"""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "hostcpp", "_bin")


def _need(name):
    exe = os.path.join(BIN, name)
    if not os.path.exists(exe):
        pytest.skip(f"{exe} not built (needs /root/reference at build time)")
    return exe


def _bench(extra, nproc):
    cmd = [sys.executable]
    if nproc > 1:
        cmd += ["-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(nproc), "--master-addr", "127.0.0.1",
                "--master-port", "29533"]
    cmd += [os.path.join(ROOT, "bench.py"), "--gpus", str(nproc), "--workload", "ensemble", "--steps", "1", "--warmup", "0"] + extra
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-3000:]
    return json.loads(p.stdout.strip().splitlines()[-1])


def test_partition_and_gather_do_not_depend_on_rank_count():
    _need("traj_dump_ref")
    args = ["--host-logic-check", "--functors", "host", "--members", "7", "--size", "150", "--threads", "2"]
    one = _bench(args, 1)
    two = _bench(args, 2)
    assert one["n_gpus"] == 1 and two["n_gpus"] == 2
    assert one["failed_members"] == 0 and two["failed_members"] == 0
    assert one["accepted_steps_total"] == two["accepted_steps_total"] > 0
    assert one["state_checksum"] == two["state_checksum"]  # members are bit-identical wherever they run
    assert two["config"]["members_per_gpu"] == 4  # contiguous blocks of ceil(7/2)
    assert two["scaling"] == "strong" and two["metric"] == "ensemble systems/s"


def test_lockstep_dealing_and_gather_do_not_depend_on_rank_count():
    """The lock-step driver's sharding (bench.py: run_ensemble_device): members dealt round-robin to the ranks, every rank's
    share integrated as the lanes of one batch, one all_gather, rank-major blocks transposed back to member order.  On CPU:
    gloo, world_size 1 and 2, the driver compiled over the emulated batch entry points (tests/emu) -- every member must
    come out bit-identical and in the same place whatever the number of ranks."""
    if not os.path.exists(os.path.join(BIN, "libensdevice_emu.so")):
        pytest.skip("hostcpp/_bin/libensdevice_emu.so not built")
    args = ["--host-logic-check", "--members", "7", "--size", "60"]
    one = _bench(args, 1)
    two = _bench(args, 2)
    assert one["n_gpus"] == 1 and two["n_gpus"] == 2
    assert one["failed_members"] == 0 and two["failed_members"] == 0
    assert one["accepted_steps_total"] == two["accepted_steps_total"] > 0
    assert one["state_checksum"] == two["state_checksum"]
    assert two["config"]["members_per_gpu"] == 4 and "round-robin" in two["config"]["parallelism"]
    assert two["lane_iterations_per_step"] == one["lane_iterations_per_step"]
    assert two["scaling"] == "strong" and two["metric"] == "ensemble systems/s"


def _members(exe, out, members, n, env=None):
    cmd = [exe, "ensemble", f"members={members}", "first=0", f"count={members}", "threads=4", f"n={n}", f"out={out}"]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900, env=env)
    assert p.returncode == 0, p.stderr[-3000:]
    info = json.loads(p.stdout.strip().splitlines()[-1])
    data = np.fromfile(out, dtype=np.float64).reshape(members, 2 * n + 2)
    os.remove(out)
    return info, data


@pytest.mark.gpu
@pytest.mark.parametrize("n", [300, 4000])
def test_ensemble_members_match_reference(tmp_path, n):
    members = 8 if n == 4000 else 12
    _, ref = _members(_need("traj_dump_ref"), str(tmp_path / "ref.f64"), members, n)
    info, got = _members(_need("traj_dump"), str(tmp_path / "gpu.f64"), members, n)
    assert info["impl"] == "b200" and info["failed"] == 0
    assert np.array_equal(got[:, 0], ref[:, 0])  # exit codes
    assert np.all(np.abs(got[:, 1] - ref[:, 1]) <= np.maximum(0.02 * ref[:, 1], 0.5))  # accepted steps +-2 %
    atol = rtol = 1e-6  # SolverOptions defaults (dae-cpp/solver-options.hpp)
    err = np.abs(got[:, 2:] - ref[:, 2:])
    assert np.all(err <= 10.0 * (atol + rtol * np.abs(ref[:, 2:])))
