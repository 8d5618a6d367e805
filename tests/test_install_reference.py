"""``niftyb200.install(nifty8)`` under the reference's OWN tests, unchanged: the transform round-trip /
normalisation tests (``test/test_operators/test_fft_operator.py``) and the adjointness / inverse / linearity
checks of ``FFTOperator`` and ``HartleyOperator`` (``test/test_operators/test_adjoint.py:149-160`` via
``ift.extra.check_linear_operator``) run in a sub-process whose ``ducc_dispatch`` seam points at this backend
(tests/ref_plugin.py).  Here: the emulator build (no GPU in the build container); with a GPU and the
reference checked out the same test runs the CUDA library.  Skipped where /root/reference is absent."""

import os
import re
import subprocess
import sys
import tempfile

import pytest

import conftest

REF = os.environ.get("NIFTYB200_REFERENCE", "/root/reference")
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "test", "test_operators")),
                                reason="reference checkout not present")


def run_reference_tests(files, keyword, backend_env):
    env = dict(os.environ, PYTHONDONTWRITEBYTECODE="1", PYTHONPATH=os.path.join(conftest.ROOT, "tests"), **backend_env)
    cmd = [sys.executable, "-m", "pytest", "-q", "-p", "no:cacheprovider", "-p", "ref_plugin", "--rootdir", REF,
           "-k", keyword] + [os.path.join(REF, "test", "test_operators", f) for f in files]
    with tempfile.TemporaryDirectory() as tmp:        # the reference's setup_function deletes *.pickle/*.png in cwd
        res = subprocess.run(cmd, cwd=tmp, env=env, capture_output=True, text=True, timeout=1500)
    tail = res.stdout[-3000:] + res.stderr[-2000:]
    assert res.returncode == 0, tail
    m = re.search(r"niftyb200-backend-launches: (\d+)", res.stdout)
    assert m and int(m.group(1)) > 100, tail          # the seam really went through this backend
    passed = re.search(r"(\d+) passed", res.stdout)
    assert passed and "failed" not in res.stdout.splitlines()[-1], tail
    return int(passed.group(1))


def test_reference_fft_operator_tests_on_emulator():
    n = run_reference_tests(["test_fft_operator.py"], "test_", {"NIFTYB200_EMU_SO": conftest.build_emulator()})
    assert n >= 100


def test_reference_adjointness_tests_on_emulator():
    n = run_reference_tests(["test_adjoint.py"], "testFFT or testHartley", {"NIFTYB200_EMU_SO": conftest.build_emulator()})
    assert n >= 16


@pytest.mark.gpu
def test_reference_fft_operator_tests_on_cuda():
    run_reference_tests(["test_fft_operator.py", "test_adjoint.py"], "test_fft or test_normalisation or testFFT or testHartley", {})
