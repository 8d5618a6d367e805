"""ctypes binding of ``libniftyb200.so`` (C ABI: ``include/niftyb200.h``).

There is no CPU fallback: if the CUDA library is missing or cannot be loaded, every
entry point of this package raises.  (The CPU-only unit tests inject the sequential
thread *emulator* build of the same sources through :func:`use_library_for_testing`;
that build lives in ``tests/emu`` and is never shipped or loaded by this package on
its own.)
"""

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_int, c_int32, c_int64, c_void_p

F32, F64 = 0, 1
# layout of the device-resident CG state (NB2_CG_* of include/niftyb200.h)
CG_GAMMA, CG_ENERGY, CG_CURV, CG_ALPHA, CG_BETA, CG_SUMS, CG_ITER, CG_INFO = 0, 1, 2, 3, 4, 5, 8, 9
CG_MINITER, CG_MAXITER, CG_RESNORM, CG_ABSDELTA, CG_EPS, CG_TINY, CG_STATE_SIZE = 10, 11, 12, 13, 14, 15, 16
NONCANONICAL, CANONICAL = 0, 1

_HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_PATH = os.path.join(_HERE, "libniftyb200.so")


class PointOp(ctypes.Structure):
    """``nb2_pointop`` (include/niftyb200.h)."""

    _fields_ = [
        ("amp_mode", c_int32),
        ("amp_flags", c_int32),
        ("amp", c_void_p),
        ("scale", c_double),
        ("mul", c_void_p),
        ("add", c_void_p),
        ("add_scalar", c_double),
        ("dot", c_void_p),
        ("dot_out", c_void_p),
        ("bin_x", c_void_p),
        ("bin_out", c_void_p),
        ("bin_fixed", c_void_p),
        ("bin_fixed_scale", c_double),
        ("pre_add", c_void_p),
        ("pre_scale", c_double),
        ("pre_out", c_void_p),
        ("pre_scale_dev", c_void_p),
    ]


class NativeError(RuntimeError):
    pass


class Library:
    """Loaded shared library with typed entry points."""

    SYMBOLS = {
        "nb2_version": (c_int, []),
        "nb2_last_error": (c_char_p, []),
        "nb2_sm_count": (c_int, []),
        "nb2_plan_create": (c_int, [c_int, POINTER(c_int64), c_int, POINTER(c_int32), c_int, POINTER(c_void_p)]),
        "nb2_plan_destroy": (c_int, [c_void_p]),
        "nb2_plan_is_fast": (c_int, [c_void_p]),
        "nb2_plan_kind": (c_int, [c_void_p]),
        "nb2_plan_is_c2c_fast": (c_int, [c_void_p]),
        "nb2_plan_set_mirror": (c_int, [c_void_p, c_int, c_int]),
        "nb2_plan_ksq_size": (c_int64, [c_void_p]),
        "nb2_hartley": (c_int, [c_void_p, c_void_p, c_void_p, c_int, POINTER(PointOp), POINTER(PointOp), c_void_p]),
        "nb2_expand_table": (c_int, [c_void_p, c_void_p, c_double, c_void_p, c_void_p]),
        "nb2_hartley_pass": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, POINTER(PointOp), POINTER(PointOp),
                                     c_void_p]),
        "nb2_strided_pass_remap": (c_int, [c_void_p, c_int, c_void_p, c_void_p, POINTER(PointOp), c_void_p, c_int64,
                                           c_void_p, c_int64, c_int64, c_int64, c_int64, c_int64, c_void_p]),
        "nb2_device_alloc": (c_int, [c_int64, POINTER(c_void_p)]),
        "nb2_device_free": (c_int, [c_void_p]),
        "nb2_ipc_export": (c_int, [c_void_p, c_void_p]),
        "nb2_ipc_open": (c_int, [c_void_p, POINTER(c_void_p)]),
        "nb2_ipc_close": (c_int, [c_void_p]),
        "nb2_strided_pass_scatter": (c_int, [c_void_p, c_int, c_void_p, c_void_p, POINTER(PointOp), c_void_p, c_void_p,
                                     c_void_p, c_void_p]),
        "nb2_launch_count": (c_int64, []),
        "nb2_tma_launch_count": (c_int64, []),
        "nb2_c2c": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_double, c_void_p]),
        "nb2_vdot": (c_int, [c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
        "nb2_vdot_noalloc": (c_int, [c_int, c_int64, c_void_p, c_void_p, c_void_p, c_void_p]),
        "nb2_axpby": (c_int, [c_int, c_int64, c_double, c_void_p, c_double, c_void_p, c_void_p]),
        "nb2_cg_update": (c_int, [c_int, c_int64, c_int, c_double, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p]),
        "nb2_amplitude_forward": (c_int, [c_int64, c_void_p, c_void_p, c_void_p, c_double, c_void_p, c_double, c_double,
                                          c_double, c_double, c_double, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                          c_int, c_void_p]),
        "nb2_amplitude_jvp": (c_int, [c_int64, c_void_p, c_void_p, c_void_p, c_double, c_void_p, c_double, c_double, c_double,
                                      c_double, c_double, c_int, c_void_p, c_void_p, c_void_p, POINTER(c_double), c_void_p,
                                      c_void_p, c_void_p]),
        "nb2_amplitude_vjp": (c_int, [c_int64, c_void_p, c_void_p, c_void_p, c_double, c_void_p, c_double, c_double, c_double,
                                      c_double, c_double, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_void_p]),
        "nb2_los_forward": (c_int, [c_int, c_int, POINTER(c_int64), POINTER(c_double), c_int64, c_int, c_void_p, c_int64, c_void_p, c_int64, c_void_p,
                                    c_void_p, c_void_p]),
        "nb2_los_adjoint": (c_int, [c_int, c_int, POINTER(c_int64), POINTER(c_double), c_int64, c_int, c_void_p, c_int64, c_void_p, c_int64, c_void_p,
                                    c_void_p, c_void_p]),
        "nb2_cg_decide": (c_int, [c_void_p, c_void_p, c_int, c_void_p]),
        "nb2_set_cg_fusion": (c_int, [c_void_p, c_void_p]),
        "nb2_set_launch_gate": (c_int, [c_void_p]),
        "nb2_cg_update_dev": (c_int, [c_int, c_int64, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_void_p]),
        "nb2_axpby_dev": (c_int, [c_int, c_int64, c_double, c_void_p, c_void_p, c_double, c_void_p, c_void_p, c_void_p]),
        "nb2_hartley_host": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_double]),
        "nb2_c2c_host": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_double]),
        "nb2_vdot_host": (c_int, [c_int, c_int64, c_void_p, c_void_p, c_void_p]),
    }

    def __init__(self, path, device_type):
        if not os.path.exists(path):
            raise NativeError(
                f"niftyb200: native library not found at {path}; build it with "
                "`python -c 'import __graft_entry__ as g; g.build()'` (there is no CPU fallback)")
        self.path = path
        self.device_type = device_type  # torch device type the pointers must live on
        self.cdll = ctypes.CDLL(path)
        for name, (res, args) in self.SYMBOLS.items():
            fn = getattr(self.cdll, name)
            fn.restype = res
            fn.argtypes = args
            setattr(self, name, fn)

    def check(self, rc, what):
        if rc != 0:
            msg = self.nb2_last_error()
            raise NativeError(f"niftyb200: {what} failed (code {rc}): {msg.decode() if msg else ''}")


_lib = None


def get():
    """Return the CUDA library, loading it on first use.  Raises if unavailable."""
    global _lib
    if _lib is None:
        import torch
        if not torch.cuda.is_available():
            raise NativeError("niftyb200: no CUDA device available (there is no CPU fallback)")
        # NIFTYB200_LIB: developer knob to load another CUDA build of the same library (tuning)
        _lib = Library(os.environ.get("NIFTYB200_LIB", DEFAULT_PATH), "cuda")
    return _lib


def use_library_for_testing(path, device_type="cpu"):
    """TESTS ONLY: route calls to another build of the same C ABI (the emulator)."""
    global _lib
    _lib = Library(path, device_type) if path is not None else None
    return _lib
