#!/usr/bin/env python
"""bench.py -- refactor+solve time per Newton iteration of the dae-cpp hot path on B200.

A "step" is ONE quasi-Newton iteration with factorisation enabled (dae-cpp/solver.hpp:431-627):
re-assemble J from its triplet values, b = M*dxdt - f, Jb = J - alpha*M, numeric sparse-LU
refactorisation, triangular solves, error checks and xk += dx.

  value   : device time per iteration with the triplet values / f already resident in HBM
            (daecpp_b200_iterations_resident: blocks of 10 iterations, iteration k+1 enqueued behind k, every
            iteration's flag words read by the host), CUDA events on the context's stream
  e2e     : the same iteration through the host-buffer C-ABI calls the overlay solver.hpp makes
            (set_jacobian[_values] -> iteration_host), H2D/D2H inside
  e2e_device_functors : C3 once more with the equations evaluated on the device (dae-cpp/device-functor.hpp,
            daecpp::solve_ensemble_device with one lane): wall clock per Newton iteration of a whole integration
  ensemble_1gpu : the N = 1 point of the C5 curve the N > 1 runs continue
  roofline: the numeric refactorisation phase against the measured HBM copy bandwidth
  cpu_baseline / --impl reference: the unmodified reference (Eigen::SparseLU behind
            oracle/_ref/libdaeref.so, else the C port) doing the same iteration on the host

Default workload: N = 1 GPU -> C3 (flame-1D n = 10^6, the config the metric is quoted on), with the other
single-GPU configs (C2, C4) and the C5 ensemble at 1 GPU in a `configs` block of the same JSON line;
N > 1 GPUs -> the C5 ensemble (4096 perovskite members partitioned across the ranks, metric
"ensemble systems/s", strong scaling): the one workload of BASELINE.json that shards.
A single sparse LU does not shard: `--workload flame1d --gpus N` runs independent replicas (weak scaling).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "refactor+solve us per Newton iter"
UNIT = "us"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto", "flame1d", "perovskite", "grid2d", "ensemble", "trajectory"],
                    help="auto: flame1d (C3) on one GPU, the C5 ensemble on several")
    ap.add_argument("--functors", default="device", choices=["device", "host"],
                    help="ensemble: device = lock-step batched lanes with device functors (dae-cpp/ensemble.hpp: solve_ensemble_device); "
                         "host = one context per member, host functors on a thread pool (solve_ensemble)")
    ap.add_argument("--no-extra-configs", action="store_true", help="C3 line without the `configs` block (C2, C4, C5 at one GPU)")
    ap.add_argument("--n", "--size", dest="n", type=int, default=0, help="flame1d: unknowns; perovskite: N (system 2N); grid2d: side")
    ap.add_argument("--problem", default="perovskite", choices=["flame", "perovskite", "flame1d", "grid2d"],
                    help="trajectory: which daecpp::solve run (C1 flame, C2 perovskite as shipped; flame1d / grid2d take --size)")
    ap.add_argument("--members", type=int, default=4096, help="ensemble: members of the whole sweep (C5)")
    ap.add_argument("--threads", type=int, default=0, help="ensemble: members in flight per GPU (0: host cores / GPUs)")
    ap.add_argument("--host-logic-check", action="store_true",
                    help="ensemble: run the partition + gather logic on CPU (gloo, reference driver) -- used by tests/")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--option", action="append", default=[], metavar="KEY=VALUE",
                    help="daecpp_b200_set_option before the first call (A/B runs of one kernel variant on one box)")
    return ap.parse_args()


# ------------------------------------------------------------------------------ workloads
def make_workload(name, size):
    """Returns dict(n, J, M, x0, f, dt, order, label, nlu_ref) -- the triplet streams the user
    functors of the BASELINE configs hand to the solver (tests/problems.py)."""
    import problems as P
    if name == "flame1d":
        n = size or 1_000_000
        x0 = P.flame1d_state(n)
        J = P.flame1d_jacobian(x0)
        M = P.identity_mass(n)
        f = P.flame1d_rhs(x0)
        nlu = 4 * n  # Eigen/COLAMD on a tridiagonal matrix: nnz(L) = nnz(U) = 2n
        label = f"C3 flame-1D n={n} tridiagonal (BASELINE configs[2])"
    elif name == "perovskite":
        N = size or 4000
        n = 2 * N
        x0 = np.concatenate([np.ones(N), np.zeros(N)])
        J = P.perovskite_jacobian(x0, N)
        M = P.perovskite_mass(N)
        f = np.linspace(-1.0, 1.0, n)
        nlu = 55_993 if N == 4000 else None
        label = f"C2 perovskite N={N} (n={n}) (BASELINE configs[1])"
    else:
        m = size or 1024
        n = m * m
        x0 = P.grid2d_state(m)
        J = P.grid2d_jacobian(x0, m)
        M = P.identity_mass(n)
        f = np.linspace(-1.0, 1.0, n)  # a non-zero right-hand side: the parity figure below compares dx with the reference's
        nlu = 153_876_742 if m == 1024 else None
        label = f"C4 grid-2D {m}x{m} 5-point (BASELINE configs[3])"
    return dict(n=n, J=J, M=M, x0=x0, f=f, dt=np.array([1e-2, 0.0, 0.0, 0.0]), order=1, label=label, nlu_ref=nlu,
                name=name)


def algorithmic_bytes(w, nnz_jb):
    """SURVEY.md 8(d) roofline table, bytes per Newton iteration and phase."""
    n = w["n"]
    nJ, nM = len(w["J"][0]), len(w["M"][0])
    order = w["order"]
    out = {
        "assembly": 12 * (nJ + nM) + 8 * nnz_jb,
        "residual": 8 * n * (order + 1) + 12 * nM + 16 * n,
        "update": 32 * n,
    }
    if w["nlu_ref"] is not None:
        out["refactor"] = 12 * nnz_jb + 12 * w["nlu_ref"]
        out["solve"] = 12 * w["nlu_ref"] + 24 * n
    return out


# ------------------------------------------------------------------------------ clocks
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.rows = []
        self.proc = None
        self.device = device

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            p = [q.strip() for q in r.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0])); smax.append(float(p[1]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if p[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------ CPU reference
def cpu_reference_iterations(w, budget_s, max_iters, warmup=0):
    """The reference's own Newton linear step on ONE host core -- the reference is single-threaded.
    Held the way the reference's solve() holds it (oracle/ref_shim.cpp: daeref_newton_open): triplet buffers, the converted
    static mass matrix and the SparseLU object persist; a timed iteration is J.convert, b = M*dxdt - f, Jb -= M*alpha,
    compute (analyzePattern + factorize, redone every time: solver.hpp:537) and solve.
    Returns (us_per_iter, iters, kind, dx, buckets_ms, per_iteration_us)."""
    import oracle_lib as O  # the checker; only this leg may touch oracle/
    n = w["n"]
    JA, Ji, Jj = [np.ascontiguousarray(a) for a in w["J"]]
    MA, Mi, Mj = [np.ascontiguousarray(a) for a in w["M"]]
    alpha = 1.0 / w["dt"][0]
    dxdt = np.zeros(n)  # xk == x[0] at the first iteration of a BDF-1 step
    f = np.ascontiguousarray(w["f"])
    dx = np.zeros(n)
    times = np.zeros(4)
    acc = np.zeros(4)
    session = None
    if O.ref is not None and hasattr(O.ref, "daeref_newton_open"):
        kind = "reference"
        session = O.ref.daeref_newton_open(n, len(JA), JA, Ji, Jj, len(MA), MA, Mi, Mj, dxdt, f)

        def one():
            rc = O.ref.daeref_newton_iterate(session, alpha, dx.ctypes.data, times.ctypes.data)
            assert rc == 0, rc
    else:
        kind = "port"

        def one():
            csc = O.assemble_jb(O.port, n, (JA, Ji, Jj), (MA, Mi, Mj), alpha)
            b = -f
            assert O.port.oracle_lu_solve(n, *csc, b, dx) == 0
    for _ in range(warmup):
        one()
    per = []
    t_start = time.perf_counter()
    while len(per) < max_iters:
        t0 = time.perf_counter()
        one()
        per.append((time.perf_counter() - t0) * 1e6)
        acc += times
        if time.perf_counter() - t_start > budget_s:
            break
    if session is not None:
        O.ref.daeref_newton_close(session)
    iters = len(per)
    return float(np.mean(per)), iters, kind, dx.copy(), (acc / iters).tolist(), per


def spread(samples):
    a = np.asarray(samples, dtype=float)
    return {"min": float(a.min()), "median": float(np.median(a)), "max": float(a.max()), "samples": int(a.size)}


def c3_config(w, nnz_jb=None):
    """The `config` object both arms print for the single-system workloads (same keys, same values)."""
    n = w["n"]
    cfg = {"workload": w["label"], "n": n, "nnz_triplets": len(w["J"][0]),
           "parallelism": "1 system per GPU (replicas only; sparse LU of one DAE does not shard)",
           "l2_policy": "working set (factors + tables + triplets + state) exceeds the 126 MB L2" if n >= 500_000
                        else "working set fits L2 (latency-bound config)"}
    return cfg


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w = make_workload(args.workload, args.n)
    # each reference iteration takes 0.45 s (C3) to 40 s (C4): the sample is bounded, the warm-up is the port arm's
    us, iters, kind, _, buckets, per = cpu_reference_iterations(w, budget_s=150.0, max_iters=args.steps, warmup=min(args.warmup, 5))
    sample = f"{iters} full-size Newton iterations of {w['label']} on 1 host core"
    line = {
        "impl": "reference", "metric": METRIC, "value": us, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": us / 1e3, "higher_is_better": False, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": c3_config(w), "spread_us": spread(per), "steps_timed": iters,
        "cpu_baseline": {"value": us, "unit": UNIT, "cores": 1, "kind": kind, "sample": sample,
                         "buckets_ms": dict(zip(["convert", "linear_algebra", "compute", "solve"], buckets))},
        "e2e": {"value": us, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------ whole trajectories (C1, C2): BDF steps/s
TRAJ_METRIC = "BDF steps/s"


def run_trajectory(args, reference):
    """One whole daecpp::solve (host functors, autodiff, observer, step control) through the overlay
    header + libdaecpp_b200.so, or -- reference arm -- through the unmodified reference headers; accepted
    BDF steps per second of wall time, host functor time reported apart."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    exe = _driver("traj_dump_ref" if reference else "traj_dump")
    kv = [f"n={args.n}"] if args.n else []
    kv.append("repeat=2")  # the second solve of the process is reported: CUDA context creation and module load are one-off
    clocks = None
    sampler = None
    if not reference:
        import torch
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
        sampler = ClockSampler(0)

    def once():
        p = subprocess.run([exe, args.problem] + kv, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        if p.returncode != 0:
            raise SystemExit(f"bench.py: {exe} {args.problem} failed ({p.returncode}): {p.stderr[-2000:]}")
        return json.loads(p.stdout.strip().splitlines()[-1])

    warm = 0 if reference else max(1, min(args.warmup, 3))
    for _ in range(warm):
        once()
    if sampler:
        sampler.start()
    runs = [once() for _ in range(1 if reference else max(1, min(args.steps, 5)))]
    if sampler:
        clocks = sampler.stop()
    r = sorted(runs, key=lambda q: q["wall_ms"])[len(runs) // 2]
    v = r["steps"] / (r["wall_ms"] * 1e-3)
    line = {"metric": TRAJ_METRIC, "value": v, "unit": "steps/s", "n_gpus": 1, "steps": len(runs), "warmup": warm,
            "ms_per_step": r["wall_ms"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"whole daecpp::solve of {args.problem} (n={r['n']}), one process, functors on the host",
                       "timing": "wall clock of the second daecpp::solve of a process (the first, with the one-off process start-up, is "
                                 "reported apart), median run; a step is one accepted BDF time step"},
            "trajectory": {"accepted_steps": r["steps"], "exit_code": r["status"], "rhs_calls": r["rhs_calls"],
                           "jac_calls": r["jac_calls"], "wall_ms": r["wall_ms"], "first_solve_of_process_ms": r["first_wall_ms"],
                           "host_functor_ms": r["ms_rhs_functor"] + r["ms_jac_functor"], "t_final": r["t_final"]},
            "e2e": {"value": v, "unit": "steps/s", "h2d_bytes_per_step": None, "d2h_bytes_per_step": None,
                    "note": "the run is the public API end to end (host buffers, functors, observer)"},
            "roofline": None}
    if reference:
        line.update({"impl": "reference", "gpu_launches": 0,
                     "cpu_baseline": {"value": v, "unit": "steps/s", "cores": 1, "kind": "reference",
                                      "sample": "one whole run of the unmodified reference (Eigen::SparseLU), 1 host core"}})
    else:
        line.update({"clocks": clocks, "gpu_launches": None})
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------ ensemble (C5)
ENS_METRIC = "ensemble systems/s"


def _driver(name):
    exe = os.path.join(ROOT, "hostcpp", "_bin", name)
    if not os.path.exists(exe):
        raise SystemExit(f"bench.py: {exe} is not built (python -c 'import __graft_entry__ as g; g.build()')")
    return exe


def _run_driver(exe, env_device, **kv):
    env = dict(os.environ)
    if env_device is not None:
        env["DAECPP_B200_DEVICE"] = str(env_device)
    cmd = [exe, "ensemble"] + [f"{k}={v}" for k, v in kv.items()]
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env)
    if p.returncode not in (0, 1):
        raise SystemExit(f"bench.py: {' '.join(cmd)} failed ({p.returncode}): {p.stderr[-2000:]}")
    return json.loads(p.stdout.strip().splitlines()[-1])


def run_ensemble_reference(args):
    """The unmodified reference headers (Eigen::SparseLU) integrating a bounded sample of the sweep on
    ALL host cores, one daecpp::solve per thread (not a single-core strawman)."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = os.cpu_count() or 1
    N = args.n or 4000
    sample = min(args.members, max(2 * cores, 8))
    vals = []
    for _ in range(max(1, min(args.steps, 3))):
        r = _run_driver(_driver("traj_dump_ref"), None, members=args.members, first=0, count=sample, threads=cores, n=N)
        vals.append(r["systems_per_s"])
    v = float(np.median(vals))
    line = {"impl": "reference", "metric": ENS_METRIC, "value": v, "unit": "systems/s", "n_gpus": args.gpus,
            "steps": len(vals), "warmup": 0, "ms_per_step": 1e3 * sample / v, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"C5 ensemble of perovskite N={N} (n={2 * N}) members, lambda log-spaced 0.5..2", "members": args.members},
            "cpu_baseline": {"value": v, "unit": "systems/s", "cores": cores, "kind": "reference",
                             "sample": f"{sample} of {args.members} members, one daecpp::solve per host thread, {cores} threads"},
            "e2e": {"value": v, "unit": "systems/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def _ensdev(emulated=False):
    """hostcpp/_bin/libensdevice.so: dae-cpp/ensemble.hpp solve_ensemble_device over the perovskite device functor
    (hostcpp/ens_device.cu, nvcc) -- called in-process so that the timed region sits between this file's barriers.
    emulated (tests only, --host-logic-check): the same driver compiled by g++ over the CPU emulation of the batch entry points."""
    import ctypes as C
    path = os.path.join(ROOT, "hostcpp", "_bin", "libensdevice_emu.so" if emulated else "libensdevice.so")
    if not os.path.exists(path):
        raise SystemExit(f"bench.py: {path} is not built (python -c 'import __graft_entry__ as g; g.build()')")
    lib = C.CDLL(path)
    lib.ensdev_perovskite.argtypes = [C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int32, C.c_double, C.c_int32, C.c_void_p, C.c_void_p]
    lib.ensdev_perovskite.restype = C.c_int
    return lib


ENS_STATS = ("rounds", "lane_iterations", "lane_refactorisations", "device_ms", "wall_ms", "failed", "accepted_steps", "checksum",
             "kernels_launched", "setup_ms", "rounds_wall_ms", "download_ms", "destroy_ms")


def ensemble_device_block(members, first, count, stride, N, device, out=None, emulated=False):
    """One lock-step integration of members first, first+stride, ... (count of them) on `device`; returns the driver's statistics."""
    os.environ["DAECPP_B200_DEVICE"] = str(device)
    st = np.zeros(len(ENS_STATS))
    _ensdev(emulated).ensdev_perovskite(members, first, count, stride, N, 10.0, 0, out.ctypes.data if out is not None else None, st.ctypes.data)
    return dict(zip(ENS_STATS, st.tolist()))


def ensemble_algorithmic_bytes(N, stats):
    """per lane iteration residual + triangular solves + update, per lane refactorisation the
    refactor bytes; C2 figures (nnz(Jb) = 39 990, nLU* = 55 993 at N = 4000), scaled with N otherwise."""
    n = 2 * N
    nnz_jb = 10 * N - 10
    nlu = 55_993 if N == 4000 else 14 * N
    nM = N
    per_iter = (8 * n * 3 + 12 * nM + 16 * n) + (12 * nlu + 24 * n) + 32 * n
    per_refac = 12 * nnz_jb + 12 * nlu
    return stats["lane_iterations"] * per_iter + stats["lane_refactorisations"] * per_refac


def run_ensemble_device(args, env=None, emit=True):
    """C5 with device functors: members dealt round-robin to the ranks, every rank integrates its share as the
    lanes of ONE batch context (one launch sequence per Newton round for all lanes), one all_gather of
    {status, steps, final state} at the end -- the only collective."""
    own_env = env is None
    emulated = bool(getattr(args, "host_logic_check", False))
    if own_env:
        env = Env(host_logic_check=emulated)
    torch, dist, world, rank, local = env.torch, env.dist, env.world, env.rank, env.local
    barrier, max_over_ranks = env.barrier, env.max_over_ranks
    dev = env.device

    def sum_over_ranks(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    N = args.n or 4000
    n = 2 * N
    members = args.members
    per = (members + world - 1) // world
    # round-robin: rank r integrates members r, r + world, ... -- every rank sees the whole lambda range, so the slow
    # members (small lambda: more steps) are spread evenly
    first, stride = rank, world
    count = len(range(rank, members, world))
    steps = max(1, min(args.steps, 20)) if args.steps_given else 5
    warm = args.warmup if emulated else max(3, min(args.warmup, 5))
    for _ in range(warm):  # whole sweeps, like the timed ones: CUDA context, module load, the context pool and the page-locked buffers come up
        ensemble_device_block(members, first, count, stride, N, local, emulated=emulated)
    out = np.zeros((count, n + 2))
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    t0 = time.perf_counter()
    acc = dict.fromkeys(ENS_STATS, 0.0)
    for _ in range(steps):
        st = ensemble_device_block(members, first, count, stride, N, local, out, emulated=emulated)
        for k in ENS_STATS:
            acc[k] += st[k]
    my_wall = time.perf_counter() - t0
    barrier()
    el = max_over_ranks(time.perf_counter() - t0)
    clocks = sampler.stop()
    dev_s = max_over_ranks(acc["device_ms"] * 1e-3)
    setup_s = max_over_ranks(acc["setup_ms"] * 1e-3)
    abytes = sum_over_ranks(ensemble_algorithmic_bytes(N, acc))
    launches = sum_over_ranks(acc["kernels_launched"])
    rounds = max_over_ranks(acc["rounds"])
    # where each rank's wall clock went, per sweep [ms]: the slowest rank sets e2e
    keys = ("setup_ms", "rounds_wall_ms", "device_ms", "download_ms", "destroy_ms", "wall_ms")
    mine_t = torch.tensor([acc[k] / steps for k in keys] + [1e3 * my_wall / steps], dtype=torch.float64, device=dev)
    if dist is not None:
        all_t = [torch.empty_like(mine_t) for _ in range(world)]
        dist.all_gather(all_t, mine_t)
    else:
        all_t = [mine_t]
    per_rank = {k: [round(float(t[i]), 2) for t in all_t] for i, k in enumerate(keys + ("call_ms",))}
    # ---- the one collective of the path
    mine = torch.from_numpy(out).to(dev)
    if count < per:
        mine = torch.cat([mine, torch.zeros(per - count, n + 2, dtype=torch.float64, device=dev)])
    gather_ms = 0.0
    if dist is not None:
        allm = torch.empty(world * per, n + 2, dtype=torch.float64, device=dev)
        env.sync()
        tg = time.perf_counter()
        dist.all_gather_into_tensor(allm, mine)
        env.sync()
        gather_ms = (time.perf_counter() - tg) * 1e3
    else:
        allm = mine
    if dist is not None:  # rank-major blocks back to member order: member m sits at block m % world, row m // world
        allm = allm.view(world, per, n + 2).transpose(0, 1).reshape(world * per, n + 2)
    allm = allm[:members]
    failed = int((allm[:, 0] != 0).sum().item())
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    peak = json.load(open(peaks_path))["hbm_gbs"] if os.path.exists(peaks_path) else 6650.0
    gbps = abytes / max(dev_s, 1e-12) / 1e9 / world  # per GPU
    line = {"metric": ENS_METRIC, "value": members * steps / max(dev_s, 1e-12), "unit": "systems/s", "n_gpus": world, "steps": steps, "warmup": warm,
            "ms_per_step": 1e3 * dev_s / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": f"C5 ensemble of perovskite N={N} (n={n}) members, lambda log-spaced 0.5..2 (BASELINE configs[4])",
                       "members": members, "members_per_gpu": per, "functors": "device (dae-cpp/device-functor.hpp: equations + dual-number shape Jacobian as kernels)",
                       "parallelism": f"members dealt round-robin to {world} GPU(s); per GPU ONE batch context, lanes advanced in lock step "
                                      "(one launch sequence + one flag read per Newton round for all lanes); no data-path collective, one final all_gather",
                       "timing": "value: CUDA-event time of all Newton rounds (state resident in HBM), max over ranks; e2e: wall clock between barriers around "
                                 "daecpp::solve_ensemble_device (context creation, pattern analysis, uploads, rounds, download of the final states), max over ranks",
                       "l2_policy": "working set (members_per_gpu x 1.6 MB) exceeds the 126 MB L2 from 80 lanes per GPU on"},
            "clocks": clocks, "failed_members": failed, "accepted_steps_total": int(allm[:, 1].sum().item()), "gather_ms": gather_ms,
            "gather_bytes": int(allm.numel() * 8), "state_checksum": float(allm[:, 2:].sum().item()),
            "rounds_per_step": rounds / steps, "lane_iterations_per_step": sum_over_ranks(acc["lane_iterations"]) / steps,
            "setup_ms_per_step": 1e3 * setup_s / steps, "per_rank_ms_per_step": per_rank,
            "e2e": {"value": members * steps / el, "unit": "systems/s", "h2d_bytes_per_step": int(world * (n * 8 + 16 * (10 * N) + 16 * N) + members * 8 + 64 * members * rounds / steps),
                    "d2h_bytes_per_step": int(members * n * 8 + 24 * members * rounds / steps),
                    "bytes_note": "h2d: per GPU the shared initial state (n doubles, replicated to the lanes on the device), the Jacobian-shape and mass "
                                  "triplets, one parameter per member, and a 64-byte lane command per member per round; d2h: the final states and a "
                                  "24-byte lane result per member per round"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "all kernels of a Newton round (functors, residual, k_chain_factor<2,1>, k_chain_forward/backward<2,1>)",
                         "achieved": gbps, "peak": peak, "unit": "GB/s", "frac": gbps / peak, "traffic": None,
                         "algorithmic_bytes": abytes / steps, "note": "per GPU; SURVEY.md 8(d) bytes x the lane iterations / refactorisations actually run"}}
    if rank == 0 and emit:
        print(json.dumps(line), flush=True)
    if own_env:
        env.close()
    return line


def run_ensemble_b200(args):
    """C5: 4096 independent perovskite solves partitioned across the ranks (contiguous blocks), every
    member a complete daecpp::solve through the overlay header; one all_gather of the final states."""
    import torch
    if not args.steps_given:
        args.steps, args.warmup = 1, min(args.warmup, 1)  # a step is the whole sweep through host functors
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    check = args.host_logic_check  # tests only: same partition / gather code, CPU tensors, reference driver
    dev = "cpu" if check else "cuda"
    if not check:
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
        torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if check:
            dist.init_process_group("gloo")
        else:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if not check:
            torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    N = args.n or 4000
    n = 2 * N
    cores = os.cpu_count() or 1
    threads = args.threads or max(1, cores // world)  # one in-flight member per host core (measured best: profiles/)
    per = (args.members + world - 1) // world
    first = rank * per
    count = max(0, min(per, args.members - first))
    exe = _driver("traj_dump_ref" if check else "traj_dump")
    out = os.path.join("/tmp", f"daecpp_ens_{os.getpid()}_{rank}.f64")
    for _ in range(args.warmup):  # a small block: brings the device, the driver and the page cache up
        _run_driver(exe, local, members=args.members, first=first, count=min(count, threads), threads=threads, n=N)
    sampler = ClockSampler(local)
    if not check:
        sampler.start()
    barrier()
    t0 = time.perf_counter()
    last = None
    for _ in range(args.steps):
        last = _run_driver(exe, local, members=args.members, first=first, count=count, threads=threads, n=N, out=out)
    barrier()
    el = time.perf_counter() - t0
    if dist is not None:
        t = torch.tensor([el], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        el = float(t.item())
    clocks = sampler.stop() if not check else None
    # ---- the one collective of the path: gather {status, steps, final state} of every member
    mine = torch.from_numpy(np.fromfile(out, dtype=np.float64).reshape(count, n + 2)).to(dev)
    os.remove(out)
    if count < per:
        mine = torch.cat([mine, torch.zeros(per - count, n + 2, dtype=torch.float64, device=dev)])
    if dist is not None:
        allm = torch.empty(world * per, n + 2, dtype=torch.float64, device=dev)
        tg = time.perf_counter()
        if check:
            parts = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(parts, mine)
            allm = torch.cat(parts)
        else:
            dist.all_gather_into_tensor(allm, mine)
            torch.cuda.synchronize()
        gather_ms = (time.perf_counter() - tg) * 1e3
    else:
        allm, gather_ms = mine, 0.0
    allm = allm[:args.members]
    failed = int((allm[:, 0] != 0).sum().item())
    steps_total = int(allm[:, 1].sum().item())
    line = {"metric": ENS_METRIC, "value": args.members * args.steps / el, "unit": "systems/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * el / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"C5 ensemble of perovskite N={N} (n={n}) members, lambda log-spaced 0.5..2 (BASELINE configs[4])",
                       "members": args.members, "members_per_gpu": per, "threads_per_gpu": threads, "host_cores": cores,
                       "parallelism": f"members partitioned in contiguous blocks over {world} GPU(s); one context + stream per "
                                      "in-flight member; no data-path collective, one final all_gather",
                       "timing": "wall clock between barriers, max over ranks (host functors are part of a solve)"},
            "clocks": clocks, "failed_members": failed, "accepted_steps_total": steps_total, "gather_ms": gather_ms,
            "gather_bytes": int(allm.numel() * 8),
            "state_checksum": float(allm[:, 2:].sum().item()),
            "rank0_driver": {k: last[k] for k in ("steps", "rhs_calls", "jac_calls", "wall_ms", "systems_per_s", "checksum")},
            "e2e": {"value": args.members * args.steps / el, "unit": "systems/s",
                    "h2d_bytes_per_step": None, "d2h_bytes_per_step": None,
                    "note": "every member is a whole daecpp::solve through host buffers: value is already end to end"},
            "gpu_launches": None, "roofline": None}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------ B200 arm
class Env:
    """Device bring-up, barrier and the max-over-ranks reduction (torch is plumbing only)."""

    def __init__(self, host_logic_check=False):
        """host_logic_check (tests only): gloo and CPU tensors, so that the dealing / gather logic around the lock-step
        driver can be exercised without a GPU (the driver itself is then the emulated one, tests/emu)."""
        import torch
        self.torch = torch
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.device = "cpu" if host_logic_check else "cuda"
        if not host_logic_check:
            if not torch.cuda.is_available():
                raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
            torch.cuda.set_device(self.local)
        self.dist = None
        if self.world > 1:
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            if host_logic_check:
                dist.init_process_group("gloo")
            else:
                dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist

    def sync(self):
        if self.device == "cuda":
            self.torch.cuda.synchronize()

    def barrier(self):
        self.sync()
        if self.dist is not None:
            self.dist.barrier()
        self.sync()

    def max_over_ranks(self, v):
        if self.dist is None:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.device)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def single_system_line(args, env, workload, size, steps, warmup, want_cpu, want_e2e):
    """One single-system workload (C2, C3 or C4) measured on this rank's GPU; returns the JSON line as a dict."""
    from gpu_lib import binding as B
    torch, dist, world, rank, local = env.torch, env.dist, env.world, env.rank, env.local
    barrier, max_over_ranks = env.barrier, env.max_over_ranks
    w = make_workload(workload, size)
    n, dt, order = w["n"], w["dt"], w["order"]
    atol = rtol = 1e-6
    lib = B.load()
    c = B.Context(n, local, lib)
    for kv in args.option:
        k, v = kv.split("=", 1)
        c.set_option(k, float(v))

    # pinned host buffers: what the overlay sparse_matrix / state vectors hand to the C ABI
    nT = len(w["J"][0])
    pA = B.PinnedArray(lib, (nT,), np.float64); pi = B.PinnedArray(lib, (nT,), np.int32); pj = B.PinnedArray(lib, (nT,), np.int32)
    pf = B.PinnedArray(lib, (n,), np.float64); px = B.PinnedArray(lib, (n,), np.float64)
    pA.array[:] = w["J"][0]; pi.array[:] = w["J"][1]; pj.array[:] = w["J"][2]; pf.array[:] = w["f"]

    # ---- cold start: pattern sort, union, symbolic analysis (once per pattern, reported apart)
    t0 = time.perf_counter()
    c.set_state(w["x0"]); c.set_mass(w["M"]); c.set_jacobian_raw(pA.array, pi.array, pj.array, nT)
    alpha = c.residual(pf.array, order, dt)
    c.refactor(alpha)
    rc, flags = c.solve_update(atol, rtol)
    cold_ms = (time.perf_counter() - t0) * 1e3
    dx_first = c.get_vector("dx")
    nnz_jb = int(c.stat("nnz_jb"))
    abytes = algorithmic_bytes(w, nnz_jb)

    def resident(k_iters, refactor=1):
        # k complete Newton iterations behind the C ABI (daecpp_b200_iterations_resident): every iteration publishes its flag
        # words and the host reads each of them, but iteration k+1 is enqueued while k runs -- where the host would have to
        # act between two iterations (failed pivot, refinement) the device cancels the queued one (api.cu)
        if k_iters > 0:
            c.iterations_resident(k_iters, refactor, order, dt, atol, rtol)

    # ---- value: K iterations, data resident in HBM
    resident(warmup)
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    barrier()
    k0 = c.stat("kernels_launched")
    # K iterations in blocks of (at most) 10, each block timed by CUDA events on the context's stream: the sum is the
    # timed region, the block means give the spread
    blocks = []
    left = steps
    while left > 0:
        kb = min(10, left)
        c.event_mark()
        resident(kb)
        blocks.append((c.event_elapsed_ms(), kb))
        left -= kb
    ms_total = sum(b[0] for b in blocks)
    barrier()
    launches = int(c.stat("kernels_launched") - k0)
    ms_total = max_over_ranks(ms_total)
    time.sleep(0.3)
    clocks = sampler.stop()
    us_iter = ms_total * 1e3 / (steps * world)

    # solve-only iterations (quasi-Newton iterations that reuse the factors)
    resident(2, 0)
    c.event_mark(); resident(steps, 0); us_solve_only = c.event_elapsed_ms() * 1e3 / steps

    # per-phase device times (events between phases; separate pass so `value` is undisturbed)
    c.set_option("phase_timing", 1)
    resident(steps)
    phases = {}
    for ph in ("assembly", "residual", "refactor", "solve", "update"):
        ms = c.stat("ms_phase_" + ph)
        phases[ph] = {"us": ms * 1e3}
        if ph == "assembly" and c.stat("jvals_from_triplets") > 0:
            phases[ph]["note"] = "fused: Jb = J - M*alpha is formed from the triplet values inside the refactorisation kernels"
            continue
        if ph == "residual" and c.stat("residual_fused") > 0:
            phases[ph]["note"] = "fused: b = M*dxdt - f is formed while the forward kernel stages its right-hand side (diagonal mass)"
            continue
        if ph == "solve" and c.stat("residual_fused") > 0:
            abytes = dict(abytes); abytes["solve"] = abytes["solve"] + abytes["residual"]
            phases[ph]["note"] = "algorithmic bytes = triangular solves + residual (fused into this phase); also holds the top-level factorisation"
        if ph == "update":
            phases[ph]["note"] = "fused: error check and xk += dx are the write-out of the last backward kernel"
            continue
        if ph in abytes and ms > 0:
            phases[ph]["algorithmic_bytes"] = abytes[ph]
            phases[ph]["GBps"] = abytes[ph] / (ms * 1e-3) / 1e9
    c.set_option("phase_timing", 0)

    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    # DRAM traffic of the refactorisation kernels from the committed `ncu --set full` capture of this
    # very command (profiles/: dram__bytes_read.sum + dram__bytes_write.sum, MB per launch), C3 only
    traffic = None
    cap = os.path.join(ROOT, "profiles", "r02_ncu_full_c3_chain_kernels.csv")  # rows: kernel, metric, value, unit (one launch)
    if workload == "flame1d" and n == 1_000_000 and os.path.exists(cap):
        import csv
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        got = {}
        for kernel, metric, value, unit in list(csv.reader(open(cap)))[1:]:
            if "k_chain_factor<1, 1>" in kernel and metric in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                got[metric] = float(value) * scale[unit]
        if len(got) == 2:
            traffic = {"bytes_per_launch": sum(got.values()), "launches_averaged": 1, "source": os.path.relpath(cap, ROOT),
                       "note": "tile-level factor kernel; the 40 MB of factor stores stay in the 126 MB write-back L2 and are re-read from "
                               "there by the forward sweep, so DRAM traffic is below the algorithmic bytes"}
    roofline = None
    if "GBps" in phases["refactor"]:
        a = phases["refactor"]["GBps"]
        roofline = {"bound": "hbm", "kernel": ("numeric refactorisation phase: k_chain_factor<B,1> (tiles; the grid-level system is factored inside the forward launch)"
                               if c.stat("chain_block") > 0 else
                               "numeric refactorisation phase (k_scatter_big + k_unit_factor + k_factor_level launches)"),
                    "achieved": a, "peak": peak, "unit": "GB/s", "frac": a / peak, "traffic": traffic,
                    "peak_source": peak_src, "algorithmic_bytes": abytes["refactor"],
                    "us": phases["refactor"]["us"]}

    if workload == "grid2d" and roofline is not None:
        # C4's refactorisation is bound by the FP64 pipe, not by HBM.  Flops = the REFERENCE structure's
        # (Eigen/COLAMD, 8.99e10 at 1024^2; extra or saved fill of this ordering is not credited); peak = a cuBLAS DGEMM
        # measured here (MEASURED_PEAKS.json holds no FP64 figure).
        ref_flops = {1024: 8.99e10}.get(size or 1024)
        a64 = torch.randn(4096, 4096, dtype=torch.float64, device="cuda"); b64 = torch.randn(4096, 4096, dtype=torch.float64, device="cuda")
        for _ in range(2):
            torch.mm(a64, b64)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(5):
            torch.mm(a64, b64)
        e1.record(); torch.cuda.synchronize()
        dgemm_tf = 5 * 2 * 4096 ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12
        del a64, b64
        own_flops = float(c.stat("factor_flops"))
        t_ref = phases["refactor"]["us"] * 1e-6
        roofline = {"bound": "fp64", "kernel": "numeric refactorisation phase (k_unit_factor + k_bf_extend/diag/rows/trsm/gemm launches)",
                    "achieved": (ref_flops / t_ref / 1e12) if ref_flops else None, "peak": dgemm_tf, "unit": "TFLOP/s",
                    "frac": (ref_flops / t_ref / 1e12 / dgemm_tf) if ref_flops else None, "traffic": None,
                    "peak_source": "cuBLAS DGEMM 4096^3 measured in this run (torch.mm, float64)",
                    "reference_flops": ref_flops, "own_structure_flops": own_flops,
                    "own_structure_TFLOPs": own_flops / t_ref / 1e12, "own_structure_frac": own_flops / t_ref / 1e12 / dgemm_tf,
                    "note": "frac credits the reference's flop count (SURVEY 8(d)); own_structure_* is the work this ordering really does",
                    "us": phases["refactor"]["us"],
                    "hbm_view": {"algorithmic_bytes": abytes["refactor"], "GBps": phases["refactor"]["GBps"], "frac_of_hbm_peak": phases["refactor"]["GBps"] / peak}}

    # ---- e2e: the host-buffer calls of the overlay header, copies inside the timed region
    e2e = None
    if want_e2e:
        # what the overlay header does per iteration (dae-cpp_b200/overlay/dae-cpp/solver.hpp): hand the
        # Jacobian over -- values only when the functor is a JacobianMatrixShape, whose (i, j) stream is
        # fixed (BASELINE configs[2] names it for C3), the full triplet stream otherwise -- then ONE
        # fused call: f upload, residual, refactor, solves, update, xk download, one synchronisation.
        fixed_pattern = workload == "flame1d"
        c.set_option("pinned_io", 1)

        def e2e_iter():
            if fixed_pattern:
                c.set_jacobian_values_raw(pA.array, nT)
            else:
                c.set_jacobian_raw(pA.array, pi.array, pj.array, nT)
            c.iteration_host(pf.array, 1, order, dt, atol, rtol, px.array)
        for _ in range(max(warmup, 3)):
            e2e_iter()
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            e2e_iter()
        torch.cuda.synchronize()
        el = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": el * 1e6 / (steps * world), "unit": UNIT,
               "h2d_bytes_per_step": int((8 if fixed_pattern else 16) * nT + 8 * n), "d2h_bytes_per_step": int(8 * n + 12),
               "path": ("set_jacobian_values + iteration_host (JacobianMatrixShape: fixed pattern)" if fixed_pattern
                        else "set_jacobian + iteration_host")}

    stats = {k: c.stat(k) for k in ("n_fronts", "n_levels", "nnz_factor", "max_front", "ms_symbolic", "lu_doubles",
                                    "cb_doubles")}
    line = {
        "metric": METRIC, "value": us_iter, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": ms_total / steps, "higher_is_better": False, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": c3_config(w), "nnz_jb": nnz_jb, "cold_start_ms": cold_ms,
        "spread_us": spread([1e3 * ms / kb for ms, kb in blocks]),
        "clocks": clocks, "gpu_launches": launches, "solve_only_us": us_solve_only, "phases": phases,
        "roofline": roofline, "e2e": e2e, "factor": stats,
        "value_protocol": "blocks of 10 iterations per daecpp_b200_iterations_resident call, CUDA events around each block; inside a block "
                          "iteration k+1 is enqueued behind k (programmatic dependent launch) while the host reads k's flag words; "
                          "--option pipeline_iterations=0 gives the one-at-a-time loop",
    }

    # ---- CPU baseline (rank 0, N = 1 only): the reference on the host, bounded sample
    if rank == 0 and world == 1 and want_cpu:
        us_cpu, iters, kind, dx_ref, buckets, _ = cpu_reference_iterations(w, budget_s=12.0, max_iters=12)
        scale = float(np.max(np.abs(dx_ref))) or 1.0
        line["cpu_baseline"] = {"value": us_cpu, "unit": UNIT, "cores": 1, "kind": kind,
                                "sample": f"{iters} full-size Newton iterations of {w['label']} on 1 host core "
                                          "(the reference is single-threaded)",
                                "buckets_ms": dict(zip(["convert", "linear_algebra", "compute", "solve"], buckets))}
        line["parity"] = {"dx_max_abs_diff_vs_reference": float(np.max(np.abs(dx_first - dx_ref))),
                          "dx_max_abs": scale}
    c.close()
    for buf in (pA, pi, pj, pf, px):
        buf.free()
    return line


def run_b200_arm(args):
    env = Env()
    line = single_system_line(args, env, args.workload, args.n, args.steps, args.warmup, not args.no_cpu_baseline, not args.no_e2e)
    if args.workload == "flame1d" and not args.n and env.world == 1 and not args.no_extra_configs:
        # the other configurations BASELINE.json names, measured on the same box in the same run (the driver records
        # this one line): C2 and C4 single systems, C5 at one GPU.  Each is the same measurement as the headline's.
        extra = {}
        for key, wl, st, wu, cpu in (("C2_perovskite", "perovskite", 200, 5, True), ("C4_grid2d_1024", "grid2d", 10, 3, False)):
            try:
                q = single_system_line(args, env, wl, 0, st, wu, cpu, True)
                extra[key] = {k: q.get(k) for k in ("metric", "value", "unit", "steps", "warmup", "spread_us", "solve_only_us", "phases", "roofline",
                                                     "e2e", "gpu_launches", "cold_start_ms", "factor", "cpu_baseline", "parity", "config")}
            except Exception as exc:  # a failure here must not take the headline down with it
                extra[key] = {"error": repr(exc)}
        if "cpu_baseline" not in extra.get("C4_grid2d_1024", {}) or extra["C4_grid2d_1024"].get("cpu_baseline") is None:
            gold = os.path.join(ROOT, "tests", "golden", "solve_c4_grid2d_1024.json")
            if os.path.exists(gold) and "error" not in extra["C4_grid2d_1024"]:
                g = json.load(open(gold))
                extra["C4_grid2d_1024"]["cpu_baseline"] = {
                    "value": None, "unit": UNIT, "cores": 1, "kind": "reference",
                    "sample": "not timed in this run: one Eigen::SparseLU compute+solve of this system takes "
                              f"{g['reference_seconds']:.0f} s on one core of the build container (tests/golden/make_golden_large.py)"}
        try:
            ens_args = argparse.Namespace(**vars(args))
            ens_args.steps, ens_args.steps_given, ens_args.warmup = 2, True, 3
            e = run_ensemble_device(ens_args, env=env, emit=False)
            extra["C5_ensemble_1gpu"] = {k: e.get(k) for k in ("metric", "value", "unit", "steps", "ms_per_step", "failed_members", "rounds_per_step",
                                                               "lane_iterations_per_step", "e2e", "gpu_launches", "roofline", "config")}
        except Exception as exc:
            extra["C5_ensemble_1gpu"] = {"error": repr(exc)}
        line["configs"] = extra
        # C3 once more, end to end with DEVICE functors (dae-cpp/device-functor.hpp): the whole flame-1D integration through
        # daecpp::solve_ensemble_device with one lane -- per Newton iteration nothing but the lane command and the flag word
        # crosses PCIe.  us per Newton iteration = wall clock of the round loop / Newton iterations run.
        try:
            import ctypes as C
            lib = _ensdev()
            lib.ensdev_flame1d.argtypes = [C.c_int32, C.c_double, C.c_int32, C.c_void_p]
            lib.ensdev_flame1d.restype = C.c_int
            os.environ["DAECPP_B200_DEVICE"] = str(env.local)
            best = None
            for _ in range(3):  # the first pass pays module load and pool warm-up
                st = np.zeros(len(ENS_STATS))
                failed = lib.ensdev_flame1d(1_000_000, 50.0, 1, st.ctypes.data)
                q = dict(zip(ENS_STATS, st.tolist()))
                if best is None or q["rounds_wall_ms"] < best["rounds_wall_ms"]:
                    best = q
            line["e2e_device_functors"] = {
                "value": 1e3 * best["rounds_wall_ms"] / best["lane_iterations"], "unit": UNIT,
                "device_us_per_iteration": 1e3 * best["device_ms"] / best["lane_iterations"],
                "newton_iterations": int(best["lane_iterations"]), "accepted_steps": int(best["accepted_steps"]), "failed": int(failed),
                "rounds": int(best["rounds"]), "setup_ms": best["setup_ms"], "final_state_download_ms": best["download_ms"],
                "whole_solve_wall_ms": best["wall_ms"], "h2d_bytes_per_step": 64, "d2h_bytes_per_step": 24,
                "path": "daecpp::solve_ensemble_device, one lane, flame-1D n=1e6 to t=50 (hostcpp/ens_device.cu: ensdev_flame1d); "
                        "equations and shape Jacobian (dual numbers) evaluated by kernels, state resident; best of 3 passes"}
        except Exception as exc:
            line["e2e_device_functors"] = {"error": repr(exc)}
        # the 1-GPU point of the 1/2/4/8-GPU ensemble curve (N > 1 runs print this metric as their `value`)
        c5 = extra["C5_ensemble_1gpu"]
        line["ensemble_1gpu"] = {k: c5.get(k) for k in ("metric", "value", "unit", "e2e", "error") if k in c5}
    if env.rank == 0:
        print(json.dumps(line), flush=True)
    env.close()


if __name__ == "__main__":
    a = parse_args()
    a.steps_given = any(x == "--steps" or x.startswith("--steps=") for x in sys.argv[1:])
    if a.workload == "auto":
        a.workload = "ensemble" if (a.gpus > 1 or int(os.environ.get("WORLD_SIZE", "1")) > 1) else "flame1d"
    if a.workload == "trajectory":
        run_trajectory(a, a.impl == "reference")
    elif a.workload == "ensemble":
        if a.impl == "reference":
            run_ensemble_reference(a)
        elif a.functors == "device":
            run_ensemble_device(a)  # (--host-logic-check: the emulated driver, gloo, CPU tensors -- tests only)
        else:
            run_ensemble_b200(a)
    elif a.impl == "reference":
        run_reference_arm(a)
    else:
        run_b200_arm(a)
