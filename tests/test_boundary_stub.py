"""The drop-in boundary exactly as INTEGRATION.md section 1 binds it: raw ctypes on the ``*_host`` entry
points (tests/integration_stub.py), against the golden vectors generated from the reference.  Runs on the
CUDA library (``-m gpu``) and, for the host logic of the same entry points, on the emulator build."""

import os

import numpy as np
import pytest

import conftest
from integration_stub import DuccDispatchStub

CONVS = ("non_canonical_hartley", "canonical_hartley")
LIB = os.path.join(conftest.PKG_DIR, "niftyb200", "libniftyb200.so")


def rel(a, b):
    return np.linalg.norm((np.asarray(a) - b).ravel()) / max(np.linalg.norm(np.asarray(b).ravel()), 1e-300)


def run_stub_against_golden(libpath, golden):
    cfg = {"hartley_convention": CONVS[0]}
    stub = DuccDispatchStub(libpath, cfg)
    g = golden["transforms"]
    n_c2c = 0
    for n in range(int(g["ncases"])):
        a = g[f"h{n}_in"]
        axes = tuple(int(x) for x in g[f"h{n}_axes"])
        axes = None if axes == (-1,) else axes
        for c in CONVS:
            cfg["hartley_convention"] = c          # read at call time, like nifty8.config (src/config.py:3-5)
            a_before = a.copy()
            out = stub.hartley(a, axes)
            assert out.dtype == np.float64 and out.shape == a.shape and np.array_equal(a, a_before)
            assert rel(out, g[f"h{n}_{c}"]) <= 1e-12, (n, c)
            out32 = stub.hartley(a.astype(np.float32), axes)
            assert out32.dtype == np.float32 and rel(out32, g[f"h{n}_{c}"]) <= 1e-5, (n, c)
        if f"h{n}_cin" in g:
            z = g[f"h{n}_cin"]
            assert rel(stub.fftn(z, axes), g[f"h{n}_fftn"]) <= 1e-12
            assert rel(stub.ifftn(z, axes), g[f"h{n}_ifftn"]) <= 1e-12
            assert rel(stub.fftn(z.astype(np.complex64), axes), g[f"h{n}_fftn"]) <= 1e-5
            n_c2c += 1
    assert n_c2c > 0
    rng = np.random.default_rng(5)
    a, b = rng.normal(size=(37, 11)), rng.normal(size=(37, 11))
    assert abs(stub.vdot(a, b) - np.vdot(a, b)) <= 1e-12 * abs(np.vdot(a, b))
    assert stub.vdot(np.arange(12), np.arange(12)) == float(np.vdot(np.arange(12), np.arange(12)))
    # errors surface as RuntimeError with the library's message (pybind11 behaviour of ducc0)
    with pytest.raises(RuntimeError, match="invalid"):
        stub._check(stub._nb2.nb2_hartley_host(None, None, None, 0, 1.0))
    with pytest.raises(RuntimeError, match="invalid"):
        stub._check(stub._nb2.nb2_c2c_host(None, None, None, 1, 1.0))
    with pytest.raises(RuntimeError, match="invalid"):
        stub._check(stub._nb2.nb2_hartley_host(stub._plan((8,), (0,), np.float64), a.ctypes.data, a.ctypes.data, 7, 1.0))
    stub.close()


def test_stub_on_emulator(golden):
    run_stub_against_golden(conftest.build_emulator(), golden)


@pytest.mark.gpu
def test_stub_on_cuda_library(golden):
    assert os.path.exists(LIB)
    run_stub_against_golden(LIB, golden)
    # a BASELINE-shaped call through the host entry point: 2048 x 2048 fp32
    stub = DuccDispatchStub(LIB)
    x = np.random.default_rng(1).standard_normal((2048, 2048), dtype=np.float32)
    y = stub.hartley(x)
    x64, y64 = x.astype(np.float64), y.astype(np.float64)
    assert abs(np.vdot(y64, y64) / (x.size * np.vdot(x64, x64)) - 1) <= 1e-5
    assert rel(stub.hartley(y) / x.size, x) <= 1e-5
    stub.close()
