"""CPU-side checks of the boundary: the C-ABI library loads, exports every symbol that
include/daecpp_b200.h declares, and refuses to work without a CUDA device (no CPU fallback)."""
import ctypes as C
import os
import re

from gpu_lib import ROOT, binding


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "daecpp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(daecpp_b200_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = binding.load()
    names = declared_symbols()
    assert len(names) >= 20
    for name in names:
        assert hasattr(lib, name), f"{name} declared in include/daecpp_b200.h but not exported"
    assert sorted(binding.SYMBOLS) == names


def test_version_string():
    lib = binding.load()
    assert b"sm_100a" in lib.daecpp_b200_version()


def test_product_library_does_not_link_the_oracle():
    # the product must never route through oracle/ (or Eigen): check the dynamic dependencies
    out = os.popen(f"ldd {binding.LIB_PATH}").read()
    assert "oracle" not in out and "daeref" not in out


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        return
    lib = binding.load()
    h = C.c_void_p()
    rc = lib.daecpp_b200_create(10, 0, C.byref(h))
    assert rc == binding.ECUDA
    assert b"no CUDA device" in lib.daecpp_b200_last_error(None)
