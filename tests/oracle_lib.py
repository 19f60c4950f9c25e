"""ctypes bindings for the two CHECKERS (test infrastructure, never the product path):

* ``port`` -- oracle/liboracle.so, the plain-C restatement (oracle/oracle.c)
* ``ref``  -- oracle/_ref/libdaeref.so, the unmodified reference headers behind a C shim
              (oracle/ref_shim.cpp); None if it was never built.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_f64 = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_i32 = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_vp = C.c_void_p


def _load(path):
    return C.CDLL(path) if os.path.exists(path) else None


port = _load(os.path.join(ROOT, "oracle", "liboracle.so"))
ref = _load(os.path.join(ROOT, "oracle", "_ref", "libdaeref.so"))

if port is not None:
    port.oracle_convert.restype = C.c_int64
    port.oracle_convert.argtypes = [C.c_int32, C.c_int64, _f64, _i32, _i32, _i32, _i32, _f64]
    port.oracle_assemble_jb.restype = C.c_int64
    port.oracle_assemble_jb.argtypes = [C.c_int32, C.c_int64, _f64, _i32, _i32, C.c_int64, _f64, _i32, _i32,
                                        C.c_double, _i32, _i32, _f64]
    port.oracle_bdf_coefficients.restype = C.c_double
    port.oracle_bdf_coefficients.argtypes = [C.c_int32, _f64, _f64]
    port.oracle_time_derivative.restype = C.c_double
    port.oracle_time_derivative.argtypes = [C.c_int32, C.c_int32, _f64, _f64, _f64, _f64]
    port.oracle_residual.restype = None
    port.oracle_residual.argtypes = [C.c_int32, _i32, _i32, _f64, _f64, _f64, _f64]
    port.oracle_error_check_update.restype = C.c_int
    port.oracle_error_check_update.argtypes = [C.c_int32, _f64, _f64, _f64, C.c_double, C.c_double, C.c_double]
    port.oracle_accept_step.restype = C.c_double
    port.oracle_accept_step.argtypes = [C.c_int32, _f64, _f64, C.c_int, C.c_double]
    port.oracle_lu_solve.restype = C.c_int
    port.oracle_lu_solve.argtypes = [C.c_int32, _i32, _i32, _f64, _f64, _f64]
    port.oracle_backward_error.restype = C.c_double
    port.oracle_backward_error.argtypes = [C.c_int32, _i32, _i32, _f64, _f64, _f64, C.POINTER(C.c_double)]

if ref is not None:
    ref.daeref_convert.restype = C.c_int64
    ref.daeref_convert.argtypes = [C.c_int32, C.c_int64, _f64, _i32, _i32, _i32, _i32, _f64, C.c_int64]
    ref.daeref_assemble_jb.restype = C.c_int64
    ref.daeref_assemble_jb.argtypes = [C.c_int32, C.c_int64, _f64, _i32, _i32, C.c_int64, _f64, _i32, _i32,
                                       C.c_double, _i32, _i32, _f64, C.c_int64]
    ref.daeref_residual.restype = None
    ref.daeref_residual.argtypes = [C.c_int32, C.c_int64, _f64, _i32, _i32, _f64, _f64, _f64]
    ref.daeref_time_derivative.restype = C.c_double
    ref.daeref_time_derivative.argtypes = [C.c_int32, C.c_int32, _f64, _f64, _f64, _f64]
    ref.daeref_sparselu.restype = C.c_int
    ref.daeref_sparselu.argtypes = [C.c_int32, _i32, _i32, _f64, _f64, _f64, C.c_int, _vp, _vp]
    ref.daeref_newton_step.restype = C.c_int
    ref.daeref_newton_step.argtypes = [C.c_int32, C.c_int64, _f64, _i32, _i32, C.c_int64, _f64, _i32, _i32,
                                       C.c_double, _f64, _f64, _f64, _vp]


    if hasattr(ref, "daeref_newton_open"):
        ref.daeref_newton_open.restype = _vp
        ref.daeref_newton_open.argtypes = [C.c_int32, C.c_int64, _f64, _i32, _i32, C.c_int64, _f64, _i32, _i32, _f64, _f64]
        ref.daeref_newton_iterate.restype = C.c_int
        ref.daeref_newton_iterate.argtypes = [_vp, C.c_double, _vp, _vp]
        ref.daeref_newton_close.restype = None
        ref.daeref_newton_close.argtypes = [_vp]


def _trip(A, i, j):
    return (np.ascontiguousarray(A, np.float64), np.ascontiguousarray(i, np.int32),
            np.ascontiguousarray(j, np.int32))


def assemble_jb(lib, n, J, M, alpha):
    """J, M = (A, i, j) triplet arrays -> (outer, inner, vals) of J - alpha*M."""
    JA, Ji, Jj = _trip(*J)
    MA, Mi, Mj = _trip(*M)
    cap = len(JA) + len(MA) + 1
    outer = np.zeros(n + 1, np.int32)
    inner = np.zeros(cap, np.int32)
    vals = np.zeros(cap, np.float64)
    if lib is port:
        nnz = lib.oracle_assemble_jb(n, len(JA), JA, Ji, Jj, len(MA), MA, Mi, Mj, alpha, outer, inner, vals)
    else:
        nnz = lib.daeref_assemble_jb(n, len(JA), JA, Ji, Jj, len(MA), MA, Mi, Mj, alpha, outer, inner, vals, cap)
    assert nnz >= 0
    return outer, inner[:nnz].copy(), vals[:nnz].copy()


def convert(lib, n, T):
    A, i, j = _trip(*T)
    cap = len(A) + 1
    outer = np.zeros(n + 1, np.int32)
    inner = np.zeros(cap, np.int32)
    vals = np.zeros(cap, np.float64)
    if lib is port:
        nnz = lib.oracle_convert(n, len(A), A, i, j, outer, inner, vals)
    else:
        nnz = lib.daeref_convert(n, len(A), A, i, j, outer, inner, vals, cap)
    return outer, inner[:nnz].copy(), vals[:nnz].copy()


def backward_error(n, csc, x, b):
    outer, inner, vals = csc
    r2 = C.c_double(0)
    eta = port.oracle_backward_error(n, outer, inner, vals, np.ascontiguousarray(x), np.ascontiguousarray(b),
                                     C.byref(r2))
    return eta, r2.value


def ref_sparselu(n, csc, b, repeat=1):
    outer, inner, vals = csc
    x = np.zeros(n)
    times = np.zeros(3)
    nnz = np.zeros(2, np.int64)
    rc = ref.daeref_sparselu(n, outer, inner, vals, np.ascontiguousarray(b), x, repeat,
                             times.ctypes.data, nnz.ctypes.data)
    return rc, x, times, nnz
