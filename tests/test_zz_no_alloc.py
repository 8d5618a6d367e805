"""``nb2_vdot_noalloc`` (include/niftyb200.h): the dot-product reduction with its scratch in a static device
array, for callers that must not allocate.  Called through the C ABI (ctypes) and compared with the
existing ``nb2_vdot`` on two calls in a row.  This file is collected LAST on purpose (
a failure here must not hide other tests) and passes alone in a fresh process as well as after
the other tests."""

import ctypes

import numpy as np
import pytest
import torch

MAX_N = 16384 * 65536        # NB2_VDOT_NOALLOC_MAX_N
F32, F64 = 0, 1


def _both(lib, a, b):
    """(nb2_vdot_noalloc, nb2_vdot) of two device tensors, through the raw C ABI."""
    code = F32 if a.dtype == torch.float32 else F64
    o1 = torch.full((), float("nan"), dtype=torch.float64, device=a.device)
    o2 = torch.full((), float("nan"), dtype=torch.float64, device=a.device)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream) if a.device.type == "cuda" else None
    pa, pb = ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(b.data_ptr())
    rc = lib.nb2_vdot_noalloc(code, a.numel(), pa, pb, ctypes.c_void_p(o1.data_ptr()), st)
    assert rc == 0, lib.nb2_last_error()
    rc = lib.nb2_vdot(code, a.numel(), pa, pb, ctypes.c_void_p(o2.data_ptr()), st)
    assert rc == 0, lib.nb2_last_error()
    return float(o1.cpu()), float(o2.cpu())


def _check_two_calls_in_a_row(lib, device, sizes):
    gen = torch.Generator(device=device).manual_seed(3)
    for dt in (torch.float64, torch.float32):
        # the second call is SMALLER than the first: partials left over from the first must not leak in
        for n in sizes:
            a = torch.randn(n, dtype=dt, device=device, generator=gen)
            b = torch.randn(n, dtype=dt, device=device, generator=gen) + 0.5 * a
            got, ref = _both(lib, a, b)
            exact = float(torch.dot(a.double(), b.double()).cpu())
            scale = float((a.double().abs() * b.double().abs()).sum().cpu())
            assert abs(got - ref) <= 1e-12 * scale, (dt, n, got, ref)         # fp64 accumulation in both
            assert abs(got - exact) <= 1e-12 * scale, (dt, n, got, exact)
    # unaligned operands (the scalar tail path of the kernel) and n = 0
    a = torch.randn(1001, dtype=torch.float32, device=device, generator=gen)
    got, ref = _both(lib, a[1:], a[:-1])
    assert abs(got - ref) <= 1e-12 * float((a[1:].double().abs() * a[:-1].double().abs()).sum().cpu())
    got, ref = _both(lib, a[:0], a[:0])
    assert got == 0.0 and ref == 0.0


def _check_rejects_oversize(lib, device):
    a = torch.zeros(8, dtype=torch.float32, device=device)
    out = torch.zeros((), dtype=torch.float64, device=device)
    n0 = lib.nb2_launch_count()
    rc = lib.nb2_vdot_noalloc(F32, MAX_N + 1, ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(a.data_ptr()),
                              ctypes.c_void_p(out.data_ptr()), None)
    assert rc == -1 and b"exceeds the maximum" in lib.nb2_last_error()
    assert lib.nb2_launch_count() == n0                    # rejected before any launch
    rc = lib.nb2_vdot_noalloc(7, 8, ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(a.data_ptr()),
                              ctypes.c_void_p(out.data_ptr()), None)
    assert rc == -1 and b"invalid argument" in lib.nb2_last_error()


@pytest.mark.gpu
def test_vdot_noalloc_matches_vdot(cuda):
    _check_two_calls_in_a_row(cuda, "cuda", (1 << 24, 100003, 5))
    _check_rejects_oversize(cuda, "cuda")


@pytest.mark.gpu
def test_vdot_noalloc_does_not_allocate(cuda):
    """Device memory in use (driver view) does not move across calls, and the call is legal during stream
    capture -- where cudaMalloc is not."""
    a = torch.randn(1 << 22, dtype=torch.float32, device="cuda")
    out = torch.zeros((), dtype=torch.float64, device="cuda")
    args = (F32, a.numel(), ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(a.data_ptr()), ctypes.c_void_p(out.data_ptr()))
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        assert cuda.nb2_vdot_noalloc(*args, ctypes.c_void_p(s.cuda_stream)) == 0      # resolves the symbol address
        s.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        for _ in range(3):
            assert cuda.nb2_vdot_noalloc(*args, ctypes.c_void_p(s.cuda_stream)) == 0
        s.synchronize()
        assert torch.cuda.mem_get_info()[0] == free0
        ref = float(out.cpu())
        out.zero_()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            rc = cuda.nb2_vdot_noalloc(*args, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, cuda.nb2_last_error()
        g.replay()
        torch.cuda.synchronize()
        assert float(out.cpu()) == ref


def test_vdot_noalloc_host_logic_on_emulator(emu):
    # same entry point on the sequential emulator build (CPU): argument checks, chunking, two calls in a row
    _check_two_calls_in_a_row(emu, "cpu", (70000, 16385, 3))
    _check_rejects_oversize(emu, "cpu")
