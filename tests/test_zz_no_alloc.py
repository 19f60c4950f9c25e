"""daecpp_b200_variability_noalloc (include/daecpp_b200.h; dae-cpp_b200/csrc/noalloc.cu): the entry point that
allocates nothing -- its scratch is a file-scope `static __device__` array -- against the existing entry point that
computes the same quantity with context-allocated scratch (daecpp_b200_accept_step, solver.hpp:703-717).
Called through the C ABI with caller-owned device buffers (torch tensors are only the device memory here), twice in
a row on different data: the second result must not see the first call's scratch contents.  Bit-exact (a maximum of
correctly rounded quotients does not depend on the reduction order), which is the bar the other variability tests use.
Collected last on purpose: the file name sorts behind every other test file."""
import ctypes as C

import numpy as np
import pytest

from gpu_lib import binding as B

pytestmark = pytest.mark.gpu


def _reference(lib, x0, xk, tol):
    """The existing entry point: a context whose x[0] = x0 and xk = xk, then accept_step with variability on."""
    c = B.Context(len(x0), 0, lib)
    c.set_state(x0)
    c.set_vector("xk", xk)
    v = c.accept_step(True, tol)
    c.close()
    return v


def _noalloc(lib, torch, x0, xk, tol, n=None):
    d_x0 = torch.from_numpy(x0).cuda()
    d_xk = torch.from_numpy(xk).cuda()
    out = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    err = C.create_string_buffer(256)
    rc = lib.daecpp_b200_variability_noalloc(d_xk.data_ptr(), d_x0.data_ptr(), len(x0) if n is None else n, tol, out.data_ptr(), None, err, 256)
    torch.cuda.synchronize()
    return rc, float(out.item()), err.value.decode()


def test_noalloc_variability_matches_accept_step_twice_in_a_row():
    import torch
    lib = B.load()
    torch.cuda.set_device(0)
    r = np.random.default_rng(12)
    tol = 1e-4
    n1 = 300_001  # 74 blocks, ragged tail
    x0a = r.standard_normal(n1)
    x0a[::7] = 1e-6           # below the tolerance: not looked at
    xka = x0a * (1.0 + 0.3 * r.standard_normal(n1))
    xka[5] = np.nan           # NaN is ignored, as in the reference
    rc, got_a, msg = _noalloc(lib, torch, x0a, xka, tol)
    assert rc == 0, msg
    want_a = _reference(lib, x0a, xka, tol)
    assert got_a == want_a and want_a > 0.5
    # second call, smaller problem AND smaller answer: stale partial maxima or a stale ticket would show
    n2 = 5_000
    x0b = 1.0 + r.random(n2)
    xkb = x0b * (1.0 + 1e-3 * r.standard_normal(n2))
    rc, got_b, msg = _noalloc(lib, torch, x0b, xkb, tol)
    assert rc == 0, msg
    want_b = _reference(lib, x0b, xkb, tol)
    assert got_b == want_b and 0.0 < want_b < 0.01
    # and the first problem again: same answer as before
    rc, got_a2, _ = _noalloc(lib, torch, x0a, xka, tol)
    assert rc == 0 and got_a2 == want_a
    # n = 0 is legal and yields 0
    rc, got_0, _ = _noalloc(lib, torch, x0b, xkb, tol, n=0)
    assert rc == 0 and got_0 == 0.0


def test_noalloc_refuses_sizes_beyond_its_fixed_scratch():
    import torch
    lib = B.load()
    torch.cuda.set_device(0)
    x = np.ones(16)
    rc, _, msg = _noalloc(lib, torch, x, x, 1e-4, n=16_777_217)  # DAECPP_B200_NOALLOC_MAX_N + 1 (nothing is touched)
    assert rc == B.EINVAL and "exceeds" in msg and "16777216" in msg
