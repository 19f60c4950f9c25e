#!/usr/bin/env python
"""bench.py -- refactor+solve time per Newton iteration of the dae-cpp hot path on B200.

A "step" is ONE quasi-Newton iteration with factorisation enabled (dae-cpp/solver.hpp:431-627):
re-assemble J from its triplet values, b = M*dxdt - f, Jb = J - alpha*M, numeric sparse-LU
refactorisation, triangular solves, error checks and xk += dx.

  value   : device time per iteration with the triplet values / f already resident in HBM
            (daecpp_b200_iteration_resident), CUDA events on the context's stream
  e2e     : the same iteration through the host-buffer C-ABI calls the overlay solver.hpp makes
            (set_jacobian -> residual -> refactor -> solve_update -> get_xk), H2D/D2H inside
  roofline: the numeric refactorisation phase against the measured HBM copy bandwidth
  cpu_baseline / --impl reference: the unmodified reference (Eigen::SparseLU behind
            oracle/_ref/libdaeref.so, else the C port) doing the same iteration on the host

N > 1 GPUs: one independent replica of the system per rank (a single sparse LU does not shard;
SURVEY.md 8(e)); value = wall time / (steps * N), weak scaling.  The ensemble workload
(--workload ensemble) partitions independent systems across ranks instead.
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
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="flame1d", choices=["flame1d", "perovskite", "grid2d"])
    ap.add_argument("--n", type=int, default=0, help="flame1d: unknowns; perovskite: N (system 2N); grid2d: side")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
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
        f = np.zeros(n)
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
    """The reference's own Newton linear step (sparse_matrix::convert x2, Jb -= M*alpha, b, Eigen
    SparseLU compute + solve) on ONE host core -- the reference is single-threaded
    (SURVEY.md section 0).  Returns (us_per_iter, iters, kind, dx, buckets_ms)."""
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
    if O.ref is not None:
        kind = "reference"

        def one():
            rc = O.ref.daeref_newton_step(n, len(JA), JA, Ji, Jj, len(MA), MA, Mi, Mj, alpha, dxdt, f, dx,
                                          times.ctypes.data)
            assert rc == 0, rc
    else:
        kind = "port"

        def one():
            csc = O.assemble_jb(O.port, n, (JA, Ji, Jj), (MA, Mi, Mj), alpha)
            b = -f
            assert O.port.oracle_lu_solve(n, *csc, b, dx) == 0
    for _ in range(warmup):
        one()
    t0 = time.perf_counter()
    iters = 0
    while iters < max_iters:
        one()
        acc += times
        iters += 1
        if time.perf_counter() - t0 > budget_s:
            break
    el = time.perf_counter() - t0
    return el / iters * 1e6, iters, kind, dx.copy(), (acc / iters).tolist()


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    w = make_workload(args.workload, args.n)
    us, iters, kind, _, buckets = cpu_reference_iterations(w, budget_s=150.0, max_iters=args.steps, warmup=min(args.warmup, 1))
    sample = f"{iters} full-size Newton iterations of {w['label']} on 1 host core"
    line = {
        "impl": "reference", "metric": METRIC, "value": us, "unit": UNIT, "n_gpus": args.gpus, "steps": iters,
        "warmup": min(args.warmup, 1), "ms_per_step": us / 1e3, "higher_is_better": False, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": w["label"], "n": w["n"], "nnz_triplets": len(w["J"][0])},
        "cpu_baseline": {"value": us, "unit": UNIT, "cores": 1, "kind": kind, "sample": sample,
                         "buckets_ms": dict(zip(["convert", "linear_algebra", "compute", "solve"], buckets))},
        "e2e": {"value": us, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------ B200 arm
def run_b200_arm(args):
    import torch  # device bring-up, barrier and the max-over-ranks reduction only
    from gpu_lib import binding as B

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    w = make_workload(args.workload, args.n)
    n, dt, order = w["n"], w["dt"], w["order"]
    atol = rtol = 1e-6
    lib = B.load()
    c = B.Context(n, local, lib)

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
        for _ in range(k_iters):
            c.iteration_resident(refactor, order, dt, atol, rtol)

    # ---- value: K iterations, data resident in HBM
    resident(args.warmup)
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    barrier()
    k0 = c.stat("kernels_launched")
    c.event_mark()
    resident(args.steps)
    ms_total = c.event_elapsed_ms()
    barrier()
    launches = int(c.stat("kernels_launched") - k0)
    ms_total = max_over_ranks(ms_total)
    time.sleep(0.3)
    clocks = sampler.stop()
    us_iter = ms_total * 1e3 / (args.steps * world)

    # solve-only iterations (quasi-Newton iterations that reuse the factors)
    resident(2, 0)
    c.event_mark(); resident(args.steps, 0); us_solve_only = c.event_elapsed_ms() * 1e3 / args.steps

    # per-phase device times (events between phases; separate pass so `value` is undisturbed)
    c.set_option("phase_timing", 1)
    resident(args.steps)
    phases = {}
    for ph in ("assembly", "residual", "refactor", "solve", "update"):
        ms = c.stat("ms_phase_" + ph)
        phases[ph] = {"us": ms * 1e3}
        if ph in abytes and ms > 0:
            phases[ph]["algorithmic_bytes"] = abytes[ph]
            phases[ph]["GBps"] = abytes[ph] / (ms * 1e-3) / 1e9
    c.set_option("phase_timing", 0)

    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    roofline = None
    if "GBps" in phases["refactor"]:
        a = phases["refactor"]["GBps"]
        roofline = {"bound": "hbm", "kernel": "numeric refactorisation phase (k_scatter + k_factor_level launches)",
                    "achieved": a, "peak": peak, "unit": "GB/s", "frac": a / peak, "traffic": None,
                    "peak_source": peak_src, "algorithmic_bytes": abytes["refactor"],
                    "us": phases["refactor"]["us"]}

    # ---- e2e: the host-buffer calls of the overlay header, copies inside the timed region
    e2e = None
    if not args.no_e2e:
        def e2e_iter():
            c.set_jacobian_raw(pA.array, pi.array, pj.array, nT)
            a = c.residual(pf.array, order, dt)
            c.refactor(a)
            c.solve_update(atol, rtol)
            lib.daecpp_b200_get_xk(c.h, px.array)
        for _ in range(max(args.warmup, 3)):
            e2e_iter()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e2e_iter()
        torch.cuda.synchronize()
        el = max_over_ranks(time.perf_counter() - t0)
        e2e = {"value": el * 1e6 / (args.steps * world), "unit": UNIT,
               "h2d_bytes_per_step": int(16 * nT + 8 * n), "d2h_bytes_per_step": int(8 * n + 8)}

    stats = {k: c.stat(k) for k in ("n_fronts", "n_levels", "nnz_factor", "max_front", "ms_symbolic", "lu_doubles",
                                    "cb_doubles")}
    line = {
        "metric": METRIC, "value": us_iter, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps, "higher_is_better": False, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": w["label"], "n": n, "nnz_triplets": nT, "nnz_jb": nnz_jb,
                   "parallelism": "1 system per GPU (replicas only; sparse LU of one DAE does not shard)",
                   "l2_policy": "working set (factors + tables + triplets + state) exceeds the 126 MB L2"
                                if n >= 500_000 else "working set fits L2 (latency-bound config)",
                   "cold_start_ms": cold_ms},
        "clocks": clocks, "gpu_launches": launches, "solve_only_us": us_solve_only, "phases": phases,
        "roofline": roofline, "e2e": e2e, "factor": stats,
    }

    # ---- CPU baseline (rank 0, N = 1 only): the reference on the host, bounded sample
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        us_cpu, iters, kind, dx_ref, buckets = cpu_reference_iterations(w, budget_s=12.0, max_iters=12)
        scale = float(np.max(np.abs(dx_ref))) or 1.0
        line["cpu_baseline"] = {"value": us_cpu, "unit": UNIT, "cores": 1, "kind": kind,
                                "sample": f"{iters} full-size Newton iterations of {w['label']} on 1 host core "
                                          "(the reference is single-threaded)",
                                "buckets_ms": dict(zip(["convert", "linear_algebra", "compute", "solve"], buckets))}
        line["parity"] = {"dx_max_abs_diff_vs_reference": float(np.max(np.abs(dx_first - dx_ref))),
                          "dx_max_abs": scale}
    c.close()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference_arm(a)
    else:
        run_b200_arm(a)
