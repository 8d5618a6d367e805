"""pytest plugin (``-p ref_plugin``) for tests/test_install_reference.py: before the reference's own test
modules are collected it imports ``nifty8`` from /root/reference (with the ``np.asfarray`` shim NumPy 2 needs,
SURVEY.md 8c) and routes its ``ducc_dispatch`` seam to this backend with ``niftyb200.install`` -- so the
reference's tests run UNCHANGED on our transforms.  At the end of the session it reports how many kernels
the backend launched (proof that the seam was really taken).  TEST INFRASTRUCTURE."""

import os
import sys

import numpy as np

if not hasattr(np, "asfarray"):
    np.asfarray = lambda a, dtype=np.float64: np.asarray(a, dtype=dtype)   # noqa: E731

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("NIFTYB200_REFERENCE", "/root/reference")
for p in (REF, os.path.join(ROOT, "joss-nifty_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    from niftyb200 import _lib
    emu = os.environ.get("NIFTYB200_EMU_SO")
    if emu:
        _lib.use_library_for_testing(emu, "cpu")
    import nifty8
    import niftyb200
    niftyb200.install(nifty8)
    from nifty8 import ducc_dispatch as dd
    from nifty8.operators import harmonic_operators as ho
    assert dd.hartley.__module__.startswith("niftyb200") and ho.fftn.__module__.startswith("niftyb200")


def pytest_sessionfinish(session, exitstatus):
    from niftyb200 import _lib
    n = int(_lib.get().nb2_launch_count())
    print(f"\nniftyb200-backend-launches: {n}")
