"""ctypes binding of libdaecpp_b200.so (include/daecpp_b200.h) used by tests/, bench.py and
__graft_entry__.py.  Pure plumbing: every call goes straight through the C ABI to the CUDA
kernels; there is no Python or CPU implementation behind it and loading fails loudly when
the library is missing."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libdaecpp_b200.so")

OK, EDECOMP, ESOLVE, EINDEX, EINVAL, ECUDA, ENOMEM = 0, 2, 3, 20, 21, 22, 23
DIVERGED, NOT_CONVERGED = 1, 2

_f64 = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_i32 = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
_vp = C.c_void_p

SYMBOLS = [
    "daecpp_b200_create", "daecpp_b200_destroy", "daecpp_b200_last_error", "daecpp_b200_version",
    "daecpp_b200_host_alloc", "daecpp_b200_host_free", "daecpp_b200_set_state", "daecpp_b200_get_xk",
    "daecpp_b200_rollback", "daecpp_b200_set_mass", "daecpp_b200_set_jacobian", "daecpp_b200_residual",
    "daecpp_b200_refactor", "daecpp_b200_solve_update", "daecpp_b200_accept_step", "daecpp_b200_set_rhs",
    "daecpp_b200_solve", "daecpp_b200_get_csc", "daecpp_b200_get_vector", "daecpp_b200_set_vector", "daecpp_b200_get_stat",
    "daecpp_b200_set_option", "daecpp_b200_iteration_resident", "daecpp_b200_iterations_resident", "daecpp_b200_event_mark", "daecpp_b200_event_elapsed_ms", "daecpp_b200_debug_unit_trace", "daecpp_b200_host_register", "daecpp_b200_host_unregister",
    "daecpp_b200_set_blocking_sync", "daecpp_b200_set_pooling", "daecpp_b200_set_jacobian_values", "daecpp_b200_iteration_host",
    "daecpp_b200_batch_create", "daecpp_b200_batch_destroy", "daecpp_b200_batch_last_error", "daecpp_b200_batch_set_pattern",
    "daecpp_b200_batch_set_mass_values", "daecpp_b200_batch_set_jacobian_values", "daecpp_b200_batch_set_rhs_values",
    "daecpp_b200_batch_set_state", "daecpp_b200_batch_get_vector", "daecpp_b200_batch_set_device_functors",
    "daecpp_b200_batch_round", "daecpp_b200_batch_get_stat", "daecpp_b200_variability_noalloc",
]

LANE_IDLE, LANE_SOLVE, LANE_REFACTOR = 0, 1, 2
LANE_KEEP, LANE_ACCEPT, LANE_ROLLBACK = 0, 1, 2


class LaneCmd(C.Structure):
    _fields_ = [("mode", C.c_int32), ("pre", C.c_int32), ("order", C.c_int32), ("period", C.c_int32), ("t", C.c_double),
                ("dt", C.c_double * 4), ("phase", C.c_int32), ("max_iterations", C.c_int32)]


class LaneResult(C.Structure):
    _fields_ = [("flags", C.c_int32), ("status", C.c_int32), ("variability", C.c_double), ("iterations", C.c_int32),
                ("reserved", C.c_int32)]


def load():
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run __graft_entry__.build() (no CPU fallback exists)")
    lib = C.CDLL(LIB_PATH)
    lib.daecpp_b200_create.argtypes = [C.c_int32, C.c_int, C.POINTER(_vp)]
    lib.daecpp_b200_destroy.argtypes = [_vp]
    lib.daecpp_b200_destroy.restype = None
    lib.daecpp_b200_last_error.argtypes = [_vp]
    lib.daecpp_b200_last_error.restype = C.c_char_p
    lib.daecpp_b200_version.restype = C.c_char_p
    lib.daecpp_b200_host_alloc.argtypes = [C.c_size_t]
    lib.daecpp_b200_host_alloc.restype = _vp
    lib.daecpp_b200_host_free.argtypes = [_vp]
    lib.daecpp_b200_host_free.restype = None
    lib.daecpp_b200_host_register.argtypes = [_vp, C.c_size_t]
    lib.daecpp_b200_host_unregister.argtypes = [_vp]
    lib.daecpp_b200_set_state.argtypes = [_vp, _f64]
    lib.daecpp_b200_get_xk.argtypes = [_vp, _f64]
    lib.daecpp_b200_rollback.argtypes = [_vp]
    lib.daecpp_b200_set_mass.argtypes = [_vp, _vp, _vp, _vp, C.c_int64]
    lib.daecpp_b200_set_jacobian.argtypes = [_vp, _vp, _vp, _vp, C.c_int64]
    lib.daecpp_b200_residual.argtypes = [_vp, _vp, C.c_int32, _f64, C.POINTER(C.c_double)]
    lib.daecpp_b200_refactor.argtypes = [_vp, C.c_double]
    lib.daecpp_b200_solve_update.argtypes = [_vp, C.c_double, C.c_double, C.c_double, C.POINTER(C.c_int32)]
    lib.daecpp_b200_accept_step.argtypes = [_vp, C.c_int32, C.c_double, C.POINTER(C.c_double)]
    lib.daecpp_b200_set_rhs.argtypes = [_vp, _f64]
    lib.daecpp_b200_solve.argtypes = [_vp, _vp]
    lib.daecpp_b200_get_csc.argtypes = [_vp, C.c_int32, _vp, _vp, _vp, C.c_int64, C.POINTER(C.c_int64)]
    lib.daecpp_b200_get_vector.argtypes = [_vp, C.c_int32, _f64]
    lib.daecpp_b200_set_vector.argtypes = [_vp, C.c_int32, _f64]
    lib.daecpp_b200_get_stat.argtypes = [_vp, C.c_char_p, C.POINTER(C.c_double)]
    lib.daecpp_b200_set_option.argtypes = [_vp, C.c_char_p, C.c_double]
    lib.daecpp_b200_iteration_resident.argtypes = [_vp, C.c_int32, C.c_int32, _f64, C.c_double, C.c_double, C.c_double,
                                                   C.POINTER(C.c_int32)]
    lib.daecpp_b200_iterations_resident.argtypes = [_vp, C.c_int32, C.c_int32, C.c_int32, _f64, C.c_double, C.c_double, C.c_double,
                                                    C.POINTER(C.c_int32)]
    lib.daecpp_b200_set_jacobian_values.argtypes = [_vp, _vp, C.c_int64]
    lib.daecpp_b200_iteration_host.argtypes = [_vp, _vp, C.c_int32, C.c_int32, _f64, C.c_double, C.c_double, C.c_double, _vp,
                                               C.POINTER(C.c_double), C.POINTER(C.c_int32)]
    lib.daecpp_b200_set_blocking_sync.argtypes = [C.c_int, C.c_int]
    lib.daecpp_b200_set_pooling.argtypes = [C.c_int]
    lib.daecpp_b200_debug_unit_trace.argtypes = [_vp, _vp, C.c_int64, C.POINTER(C.c_int64)]
    lib.daecpp_b200_event_mark.argtypes = [_vp]
    lib.daecpp_b200_event_elapsed_ms.argtypes = [_vp, C.POINTER(C.c_double)]
    lib.daecpp_b200_batch_create.argtypes = [C.c_int32, C.c_int32, C.c_int, C.POINTER(_vp)]
    lib.daecpp_b200_batch_destroy.argtypes = [_vp]
    lib.daecpp_b200_batch_destroy.restype = None
    lib.daecpp_b200_batch_last_error.argtypes = [_vp]
    lib.daecpp_b200_batch_last_error.restype = C.c_char_p
    lib.daecpp_b200_batch_set_pattern.argtypes = [_vp, _vp, _vp, C.c_int64, _vp, _vp, C.c_int64]
    lib.daecpp_b200_batch_set_mass_values.argtypes = [_vp, _vp, C.c_int32, C.c_int32]
    lib.daecpp_b200_batch_set_jacobian_values.argtypes = [_vp, _vp, C.c_int32, C.c_int32]
    lib.daecpp_b200_batch_set_rhs_values.argtypes = [_vp, _vp, C.c_int32, C.c_int32]
    lib.daecpp_b200_batch_set_state.argtypes = [_vp, _vp, C.c_int32, C.c_int32]
    lib.daecpp_b200_batch_get_vector.argtypes = [_vp, C.c_int32, _vp, C.c_int32, C.c_int32]
    lib.daecpp_b200_batch_set_device_functors.argtypes = [_vp, _vp]
    lib.daecpp_b200_batch_round.argtypes = [_vp, C.POINTER(LaneCmd), C.c_double, C.c_double, C.c_double, C.c_int32, C.c_double,
                                            C.POINTER(LaneResult)]
    lib.daecpp_b200_batch_get_stat.argtypes = [_vp, C.c_char_p, C.POINTER(C.c_double)]
    lib.daecpp_b200_variability_noalloc.argtypes = [_vp, _vp, C.c_int64, C.c_double, _vp, _vp, C.c_char_p, C.c_size_t]
    if os.environ.get("DAECPP_B200_TEST_POOLING"):  # tests: every context of the process comes out of / goes back to the pool
        lib.daecpp_b200_set_pooling(1)
    return lib


class Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"daecpp_b200 error {code}: {msg}")
        self.code = code


def _ptr(a):
    return a.ctypes.data if a is not None and a.size else None


class PinnedArray:
    """numpy view over memory from daecpp_b200_host_alloc (what the overlay sparse_matrix uses)."""

    def __init__(self, lib, shape, dtype):
        self.lib = lib
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        self.ptr = lib.daecpp_b200_host_alloc(max(n, 1))
        if not self.ptr:
            raise Error(ECUDA, "pinned allocation failed")
        buf = (C.c_char * max(n, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            self.lib.daecpp_b200_host_free(self.ptr)
            self.ptr = None


class Context:
    """One DAE system on one GPU."""

    def __init__(self, n, device=0, lib=None):
        self.lib = lib or load()
        self.n = int(n)
        h = _vp()
        rc = self.lib.daecpp_b200_create(self.n, device, C.byref(h))
        if rc:
            raise Error(rc, (self.lib.daecpp_b200_last_error(None) or b"").decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.daecpp_b200_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc, allow=()):
        if rc and rc not in allow:
            raise Error(rc, (self.lib.daecpp_b200_last_error(self.h) or b"").decode())
        return rc

    @staticmethod
    def _trip(T):
        A, i, j = T
        return (np.ascontiguousarray(A, np.float64), np.ascontiguousarray(i, np.int32),
                np.ascontiguousarray(j, np.int32))

    def set_state(self, x0):
        self._ck(self.lib.daecpp_b200_set_state(self.h, np.ascontiguousarray(x0, np.float64)))

    def get_xk(self):
        out = np.empty(self.n)
        self._ck(self.lib.daecpp_b200_get_xk(self.h, out))
        return out

    def rollback(self):
        self._ck(self.lib.daecpp_b200_rollback(self.h))

    def set_mass(self, T, allow=()):
        A, i, j = self._trip(T)
        return self._ck(self.lib.daecpp_b200_set_mass(self.h, _ptr(A), _ptr(i), _ptr(j), len(A)), allow)

    def set_jacobian(self, T, allow=()):
        A, i, j = self._trip(T)
        return self._ck(self.lib.daecpp_b200_set_jacobian(self.h, _ptr(A), _ptr(i), _ptr(j), len(A)), allow)

    def set_jacobian_raw(self, A, i, j, nT):
        """Arrays already contiguous (possibly pinned): no copies on the Python side."""
        return self._ck(self.lib.daecpp_b200_set_jacobian(self.h, A.ctypes.data, i.ctypes.data, j.ctypes.data, nT))

    def residual(self, f, order, dt4):
        f = np.ascontiguousarray(f, np.float64)
        alpha = C.c_double(0)
        self._ck(self.lib.daecpp_b200_residual(self.h, f.ctypes.data, order, np.ascontiguousarray(dt4, np.float64),
                                               C.byref(alpha)))
        return alpha.value

    def refactor(self, alpha, allow=()):
        return self._ck(self.lib.daecpp_b200_refactor(self.h, float(alpha)), allow)

    def solve_update(self, atol, rtol, max_err_abs=1e100, allow=()):
        flags = C.c_int32(0)
        rc = self._ck(self.lib.daecpp_b200_solve_update(self.h, atol, rtol, max_err_abs, C.byref(flags)), allow)
        return rc, flags.value

    def iteration_resident(self, do_refactor, order, dt4, atol, rtol, max_err_abs=1e100, allow=()):
        flags = C.c_int32(0)
        rc = self._ck(self.lib.daecpp_b200_iteration_resident(self.h, int(do_refactor), order,
                                                              np.ascontiguousarray(dt4, np.float64), atol, rtol,
                                                              max_err_abs, C.byref(flags)), allow)
        return rc, flags.value

    def iterations_resident(self, count, do_refactor, order, dt4, atol, rtol, max_err_abs=1e100, allow=()):
        flags = C.c_int32(0)
        rc = self._ck(self.lib.daecpp_b200_iterations_resident(self.h, int(count), int(do_refactor), order,
                                                               np.ascontiguousarray(dt4, np.float64), atol, rtol,
                                                               max_err_abs, C.byref(flags)), allow)
        return rc, flags.value

    def set_jacobian_values_raw(self, A, nT):
        return self._ck(self.lib.daecpp_b200_set_jacobian_values(self.h, A.ctypes.data, nT))

    def iteration_host(self, f, do_refactor, order, dt4, atol, rtol, xk_out, max_err_abs=1e100, allow=()):
        """f and xk_out: contiguous float64 arrays (pinned when option pinned_io is set)."""
        flags = C.c_int32(0)
        alpha = C.c_double(0)
        rc = self._ck(self.lib.daecpp_b200_iteration_host(self.h, f.ctypes.data, int(do_refactor), order,
                                                          np.ascontiguousarray(dt4, np.float64), atol, rtol, max_err_abs,
                                                          xk_out.ctypes.data if xk_out is not None else None,
                                                          C.byref(alpha), C.byref(flags)), allow)
        return rc, flags.value, alpha.value

    def unit_trace(self):
        cnt = C.c_int64(0)
        self._ck(self.lib.daecpp_b200_debug_unit_trace(self.h, None, 0, C.byref(cnt)))
        out = np.zeros(max(cnt.value, 1), np.int64)
        self._ck(self.lib.daecpp_b200_debug_unit_trace(self.h, out.ctypes.data, cnt.value, C.byref(cnt)))
        return out[:cnt.value - cnt.value % 4].reshape(-1, 4)

    def event_mark(self):
        self._ck(self.lib.daecpp_b200_event_mark(self.h))

    def event_elapsed_ms(self):
        ms = C.c_double(0)
        self._ck(self.lib.daecpp_b200_event_elapsed_ms(self.h, C.byref(ms)))
        return ms.value

    def accept_step(self, variability_on=True, tol=1e-4):
        v = C.c_double(0)
        self._ck(self.lib.daecpp_b200_accept_step(self.h, int(variability_on), tol, C.byref(v)))
        return v.value

    def set_rhs(self, b):
        self._ck(self.lib.daecpp_b200_set_rhs(self.h, np.ascontiguousarray(b, np.float64)))

    def solve(self, download=True):
        out = np.empty(self.n) if download else None
        self._ck(self.lib.daecpp_b200_solve(self.h, _ptr(out) if download else None))
        return out

    def get_csc(self, which):
        nnz = C.c_int64(0)
        outer = np.zeros(self.n + 1, np.int32)
        self._ck(self.lib.daecpp_b200_get_csc(self.h, which, outer.ctypes.data, None, None, 0, C.byref(nnz)))
        inner = np.zeros(max(nnz.value, 1), np.int32)
        vals = np.zeros(max(nnz.value, 1), np.float64)
        self._ck(self.lib.daecpp_b200_get_csc(self.h, which, outer.ctypes.data, inner.ctypes.data, vals.ctypes.data,
                                              nnz.value, C.byref(nnz)))
        return outer, inner[:nnz.value].copy(), vals[:nnz.value].copy()

    def get_vector(self, which):
        names = {"b": 0, "dx": 1, "dxdt": 2, "xk": 3, "x0": 4, "x1": 5, "x2": 6, "x3": 7}
        out = np.empty(self.n)
        self._ck(self.lib.daecpp_b200_get_vector(self.h, names.get(which, which), out))
        return out

    def set_vector(self, which, values):
        names = {"b": 0, "dx": 1, "xk": 3, "x0": 4, "x1": 5, "x2": 6, "x3": 7}
        self._ck(self.lib.daecpp_b200_set_vector(self.h, names.get(which, which),
                                                 np.ascontiguousarray(values, np.float64)))

    def stat(self, key):
        v = C.c_double(0)
        self._ck(self.lib.daecpp_b200_get_stat(self.h, key.encode(), C.byref(v)))
        return v.value

    def set_option(self, key, value):
        self._ck(self.lib.daecpp_b200_set_option(self.h, key.encode(), float(value)))


class Batch:
    """`lanes` systems of one sparsity pattern advanced in lock step (include/daecpp_b200.h: daecpp_b200_batch_*).
    Host-functor mode only from Python: right-hand sides and Jacobian values are uploaded as lane-major slabs."""

    VEC = {"b": 0, "dx": 1, "xk": 3, "x0": 4, "x1": 5, "x2": 6, "x3": 7}

    def __init__(self, n, lanes, device=0, lib=None):
        self.lib = lib or load()
        self.n, self.lanes = int(n), int(lanes)
        h = _vp()
        rc = self.lib.daecpp_b200_batch_create(self.n, self.lanes, device, C.byref(h))
        if rc:
            raise Error(rc, (self.lib.daecpp_b200_batch_last_error(None) or b"").decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            self.lib.daecpp_b200_batch_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc, allow=()):
        if rc and rc not in allow:
            raise Error(rc, (self.lib.daecpp_b200_batch_last_error(self.h) or b"").decode())
        return rc

    def set_pattern(self, Ji, Jj, Mi, Mj, allow=()):
        Ji, Jj, Mi, Mj = [np.ascontiguousarray(a, np.int32) for a in (Ji, Jj, Mi, Mj)]
        self.nTJ, self.nTM = len(Ji), len(Mi)
        return self._ck(self.lib.daecpp_b200_batch_set_pattern(self.h, _ptr(Ji), _ptr(Jj), len(Ji), _ptr(Mi), _ptr(Mj), len(Mi)), allow)

    def _slab(self, a, width, first, count):
        a = np.ascontiguousarray(a, np.float64)
        assert a.size == max(count, 1) * width or (count == 0 and a.size == width), (a.shape, width, count)
        return a

    def set_mass_values(self, A, first=0, count=0):
        A = self._slab(A, self.nTM, first, count)
        self._ck(self.lib.daecpp_b200_batch_set_mass_values(self.h, _ptr(A), first, count))

    def set_jacobian_values(self, A, first=0, count=None):
        count = self.lanes if count is None else count
        A = self._slab(A, self.nTJ, first, count)
        self._ck(self.lib.daecpp_b200_batch_set_jacobian_values(self.h, _ptr(A), first, count))

    def set_rhs_values(self, f, first=0, count=None):
        count = self.lanes if count is None else count
        f = self._slab(f, self.n, first, count)
        self._ck(self.lib.daecpp_b200_batch_set_rhs_values(self.h, _ptr(f), first, count))

    def set_state(self, x0, first=0, count=None):
        count = self.lanes if count is None else count
        x0 = self._slab(x0, self.n, first, count)
        self._ck(self.lib.daecpp_b200_batch_set_state(self.h, _ptr(x0), first, count))

    def get_vector(self, which, first=0, count=None):
        count = self.lanes if count is None else count
        out = np.empty((count, self.n))
        self._ck(self.lib.daecpp_b200_batch_get_vector(self.h, self.VEC.get(which, which), out.ctypes.data, first, count))
        return out

    def round(self, cmds, atol, rtol, max_err_abs=1e100, variability_on=False, var_tol=1e-4, allow=()):
        """cmds: list of dict(mode, pre, order, t, dt) per lane.  Returns (rc, [(flags, status, variability)])."""
        arr = (LaneCmd * self.lanes)()
        for k, q in enumerate(cmds):
            arr[k].mode = q.get("mode", LANE_IDLE); arr[k].pre = q.get("pre", LANE_KEEP); arr[k].order = q.get("order", 1)
            arr[k].t = q.get("t", 0.0)
            arr[k].period = q.get("period", 0); arr[k].phase = q.get("phase", 0); arr[k].max_iterations = q.get("max_iterations", 1)
            for i, v in enumerate(q.get("dt", (0.0, 0.0, 0.0, 0.0))):
                arr[k].dt[i] = v
        res = (LaneResult * self.lanes)()
        rc = self._ck(self.lib.daecpp_b200_batch_round(self.h, arr, atol, rtol, max_err_abs, int(variability_on), var_tol, res), allow)
        self.last_iterations = [r.iterations for r in res]
        return rc, [(r.flags, r.status, r.variability) for r in res]

    def stat(self, key):
        v = C.c_double(0)
        self._ck(self.lib.daecpp_b200_batch_get_stat(self.h, key.encode(), C.byref(v)))
        return v.value
