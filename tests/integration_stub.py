"""The binding INTEGRATION.md section 1 shows, as an importable module: what a NIFTy maintainer would put
into ``src/ducc_dispatch.py`` (replacing ``:61-92``) to run the ``ducc_dispatch`` seam on ``libniftyb200.so``.

Raw ``ctypes`` + NumPy only -- no torch, no ``niftyb200`` package -- through the ``*_host`` entry points of
``include/niftyb200.h``.  ``tests/test_boundary_stub.py`` runs it against the golden vectors of the reference.
"""

import ctypes

import numpy as np


class DuccDispatchStub:
    """``fftn`` / ``ifftn`` / ``hartley`` / ``vdot`` with the signatures of ``src/ducc_dispatch.py:66-86``."""

    def __init__(self, libpath, config=None):
        self._config = {"hartley_convention": "non_canonical_hartley"} if config is None else config
        nb2 = self._nb2 = ctypes.CDLL(libpath)
        nb2.nb2_plan_create.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_int64), ctypes.c_int,
                                        ctypes.POINTER(ctypes.c_int32), ctypes.c_int,
                                        ctypes.POINTER(ctypes.c_void_p)]
        nb2.nb2_plan_destroy.argtypes = [ctypes.c_void_p]
        nb2.nb2_hartley_host.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                         ctypes.c_int, ctypes.c_double]
        nb2.nb2_c2c_host.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                     ctypes.c_int, ctypes.c_double]
        nb2.nb2_vdot_host.argtypes = [ctypes.c_int, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p,
                                      ctypes.POINTER(ctypes.c_double)]
        nb2.nb2_last_error.restype = ctypes.c_char_p
        self._plans = {}

    def _check(self, rc):
        if rc:
            raise RuntimeError(self._nb2.nb2_last_error().decode())

    def _plan(self, shape, axes, dtype):
        key = (shape, axes, dtype)
        if key not in self._plans:
            p = ctypes.c_void_p()
            sh = (ctypes.c_int64 * len(shape))(*shape)
            ax = (ctypes.c_int32 * len(axes))(*axes)
            self._check(self._nb2.nb2_plan_create(len(shape), sh, len(axes), ax,
                                                  0 if dtype == np.float32 else 1, ctypes.byref(p)))
            self._plans[key] = p
        return self._plans[key]

    @staticmethod
    def _axes(a, axes):
        return tuple(range(a.ndim)) if axes is None else tuple(int(x) for x in np.atleast_1d(axes) % a.ndim)

    def hartley(self, a, axes=None):                       # replaces :75-78
        a = np.ascontiguousarray(a)
        if a.dtype != np.float32:
            a = a.astype(np.float64, copy=False)
        out = np.empty_like(a)
        conv = 1 if self._config["hartley_convention"] == "canonical_hartley" else 0
        self._check(self._nb2.nb2_hartley_host(self._plan(a.shape, self._axes(a, axes), a.dtype.type),
                                               a.ctypes.data, out.ctypes.data, conv, 1.0))
        return out

    def _c2c(self, a, axes, forward):
        a = np.asarray(a)
        cdt = np.complex64 if a.dtype in (np.float32, np.complex64) else np.complex128
        a = np.ascontiguousarray(a, dtype=cdt)
        axes = self._axes(a, axes)
        out = np.empty_like(a)
        scale = 1.0 if forward else 1.0 / float(np.prod([a.shape[i] for i in axes]))
        rdt = np.float32 if cdt == np.complex64 else np.float64
        self._check(self._nb2.nb2_c2c_host(self._plan(a.shape, axes, rdt), a.ctypes.data, out.ctypes.data,
                                           1 if forward else 0, scale))
        return out

    def fftn(self, a, axes=None):                          # replaces :66-67
        return self._c2c(a, axes, True)

    def ifftn(self, a, axes=None):                         # replaces :70-72
        return self._c2c(a, axes, False)

    def vdot(self, a, b):                                  # replaces :81-86 (real operands)
        a, b = np.asarray(a), np.asarray(b)
        if np.issubdtype(a.dtype, np.integer):
            a = a.astype(np.float64)
        if np.issubdtype(b.dtype, np.integer):
            b = b.astype(np.float64)
        dt = np.float32 if (a.dtype == np.float32 and b.dtype == np.float32) else np.float64
        a, b = np.ascontiguousarray(a, dtype=dt), np.ascontiguousarray(b, dtype=dt)
        out = ctypes.c_double()
        self._check(self._nb2.nb2_vdot_host(0 if dt == np.float32 else 1, a.size, a.ctypes.data, b.ctypes.data,
                                            ctypes.byref(out)))
        return out.value

    def close(self):
        for p in self._plans.values():
            self._nb2.nb2_plan_destroy(p)
        self._plans = {}
