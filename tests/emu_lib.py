"""Builds and loads the scalar CPU walk over the product's symbolic schedule (tests/emu/mf_emulate.cpp).
Test infrastructure: validates ordering / unit partition / blobs / slot maps without a GPU."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_SO = os.path.join(ROOT, "tests", "emu", "libmfemu.so")
_SRC = [os.path.join(ROOT, "tests", "emu", "mf_emulate.cpp"), os.path.join(ROOT, "dae-cpp_b200", "csrc", "symbolic.cpp")]
_HDR = [os.path.join(ROOT, "dae-cpp_b200", "csrc", h) for h in ("symbolic.hpp", "schedule.hpp", "unit_ops.hpp")]


def load():
    newest = max(os.path.getmtime(p) for p in _SRC + _HDR)
    if not os.path.exists(_SO) or os.path.getmtime(_SO) < newest:
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", *_SRC, "-o", _SO])
    lib = C.CDLL(_SO)
    f64 = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
    i32 = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
    lib.mfemu_solve.argtypes = [C.c_int32, i32, i32, f64, f64, f64, C.c_int32, C.c_int32, f64]
    lib.mfemu_solve.restype = C.c_int
    return lib


STAT_NAMES = ["n_fronts", "tree_height", "nnz_factor", "flops", "max_front", "lu_doubles", "cb_doubles",
              "max_pivot_block", "n_units", "n_big", "n_groups", "factor_smem_bytes"]


def solve(lib, n, csc, b, leaf_size=0, unit_smem_bytes=-1):
    outer, inner, vals = csc
    x = np.zeros(n)
    st = np.zeros(12)
    rc = lib.mfemu_solve(n, outer, inner, np.ascontiguousarray(vals), np.ascontiguousarray(b), x, leaf_size,
                         unit_smem_bytes, st)
    return rc, x, dict(zip(STAT_NAMES, st))
