import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_DIR = os.path.join(ROOT, "joss-nifty_b200")
for p in (ROOT, PKG_DIR, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")
CSRC = os.path.join(PKG_DIR, "csrc")
EMU_SO = os.path.join(ROOT, "tests", "emu", "libniftyb200_emu.so")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run with -m gpu on a B200)")


def pytest_collection_modifyitems(config, items):
    # GPU tests are skipped on boxes without a device so that a plain `pytest tests/`
    # stays green on CPU.
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def build_emulator():
    """Compile the sequential thread emulator of the kernels (tests/emu/README.md)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(ROOT, "include", "niftyb200.h")]
    if os.path.exists(EMU_SO) and all(os.path.getmtime(EMU_SO) >= os.path.getmtime(s) for s in srcs):
        return EMU_SO
    cus = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))
    cmd = ["g++", "-O1", "-std=c++17", "-x", "c++", "-DNB2_EMU", "-shared", "-fPIC", "-o", EMU_SO] + cus
    subprocess.run(cmd, check=True)
    return EMU_SO


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    class _G:
        def __getitem__(self, name):
            return np.load(os.path.join(GOLDEN, name + ".npz"))
    return _G()


@pytest.fixture()
def emu():
    """Route niftyb200 to the emulator build for one test (CPU tensors)."""
    from niftyb200 import _lib, transforms
    lib = _lib.use_library_for_testing(build_emulator(), "cpu")
    transforms._plan_cache.clear()
    yield lib
    transforms._plan_cache.clear()
    _lib.use_library_for_testing(None)


@pytest.fixture()
def cuda():
    """The product path: real CUDA library on cuda:0 (fails loudly if it is missing)."""
    from niftyb200 import _lib, transforms
    _lib.use_library_for_testing(None)
    transforms._plan_cache.clear()
    lib = _lib.get()
    assert lib.device_type == "cuda"
    yield lib
    transforms._plan_cache.clear()
