"""CPU-only: the C-ABI library loads and exports every symbol ``include/niftyb200.h`` declares
(no compute calls without a GPU), and the product path refuses to run without CUDA."""

import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "niftyb200.h")
LIB = os.path.join(ROOT, "joss-nifty_b200", "niftyb200", "libniftyb200.so")


def declared_symbols():
    txt = re.sub(r"/\*.*?\*/", "", open(HEADER).read(), flags=re.S)
    return sorted(set(re.findall(r"\b(nb2_[a-z0-9_]+)\s*\(", txt)))


def test_header_declares_expected_entry_points():
    syms = declared_symbols()
    for s in ("nb2_plan_create", "nb2_plan_destroy", "nb2_hartley", "nb2_c2c", "nb2_vdot", "nb2_axpby",
              "nb2_cg_update", "nb2_last_error"):
        assert s in syms


def test_library_exports_every_declared_symbol():
    if not os.path.exists(LIB):
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(LIB)
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} declared in include/niftyb200.h but not exported"
    lib.nb2_version.restype = ctypes.c_int
    assert lib.nb2_version() >= 100


def test_python_binding_covers_header():
    from niftyb200._lib import Library
    assert sorted(Library.SYMBOLS) == declared_symbols()


def test_no_cpu_fallback():
    import torch
    from niftyb200 import _lib
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    _lib.use_library_for_testing(None)
    import numpy as np
    import niftyb200 as nb
    with pytest.raises(_lib.NativeError):
        nb.hartley(np.ones(8))
