"""Where the first solve of a process spends its time: library load, context creation, first calls."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import importlib
import numpy as np
t = [time.perf_counter()]
def lap(name):
    t.append(time.perf_counter()); print(f"{name:32s} {1e3 * (t[-1] - t[-2]):9.2f} ms", flush=True)
B = importlib.import_module("dae-cpp_b200.binding"); import problems as P
lap("imports")
lib = B.load(); lap("dlopen libdaecpp_b200.so")
N = 4000; n = 2 * N
c = B.Context(n, 0, lib); lap("daecpp_b200_create (CUDA init)")
x = P.perovskite_state(N); c.set_state(x); c.set_mass(P.perovskite_mass(N)); lap("set_state + set_mass")
c.set_jacobian(P.perovskite_jacobian(x, N)); lap("set_jacobian (cold assembly)")
dt4 = np.array([1e-3, 0, 0, 0.0]); f = np.zeros(n)
alpha = c.residual(f, 1, dt4); lap("residual")
c.iteration_resident(1, 1, dt4, 1e-6, 1e-6); lap("first iteration (symbolic + modules)")
c.iteration_resident(1, 1, dt4, 1e-6, 1e-6); lap("second iteration")
c.close(); lap("close")
c = B.Context(n, 0, lib); lap("second create")
