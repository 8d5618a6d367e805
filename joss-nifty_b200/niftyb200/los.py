"""Line-of-sight response and the tomography metric.

``SamplingCartesianGridLOS`` mirrors ``/root/reference/src/re/extra/sampling_los.py:31-91`` (``interpolation_order=1``)
on device tensors; :class:`LOSMetric` is the metric of ``demos/1_tomography.py``: a correlated field, a point-wise
link, the LOS response and a Gaussian likelihood on the line integrals,

    M t = A H f'(s) R^T N^-1 R f'(s) H A t + t      (xi block; likelihood.py:568-576, evi.py:108-109).
"""

import ctypes

import numpy as np
import torch

from . import _lib
from .transforms import _TORCH_REAL, PointOp, _ptr, _stream, device as _device


class SamplingCartesianGridLOS:
    """``SamplingCartesianGridLOS(start, end, shape=, distances=, n_sampling_points=500)`` (:31-91)."""

    def __init__(self, start, end, *, shape, distances, n_sampling_points=500, interpolation_order=1, dtype=torch.float64):
        if interpolation_order != 1:
            raise NotImplementedError("interpolation_order=1 (multilinear) is on the GPU path")
        self.lib = _lib.get()
        self.shape = tuple(int(s) for s in shape)
        if not 1 <= len(self.shape) <= 3:
            raise ValueError("LOS response: 1 to 3 dimensions")
        nd = len(self.shape)
        self.distances = tuple(float(d) for d in np.broadcast_to(np.asarray(distances, dtype=np.float64), (nd,)))
        dev = _device()
        self.start = torch.as_tensor(np.asarray(start, dtype=np.float64)).to(dev).contiguous()
        self.end = torch.as_tensor(np.asarray(end, dtype=np.float64)).to(dev).contiguous()
        for a in (self.start, self.end):
            if a.shape[-1] != nd or a.dim() > 2:
                raise ValueError("start / end: shape (n_points, n_dim) or (n_dim,)")
        # in_axes of the reference's vmap (:84-90): a single point is shared by all lines
        self.n_los = int(max(self.start.shape[0] if self.start.dim() == 2 else 1, self.end.shape[0] if self.end.dim() == 2 else 1))
        self.ss = nd if (self.start.dim() == 2 and self.start.shape[0] > 1) else 0
        self.es = nd if (self.end.dim() == 2 and self.end.shape[0] > 1) else 0
        self.n, self.dtype = int(n_sampling_points), dtype
        self.target_shape = (self.n_los,)
        self._shape_c = (ctypes.c_int64 * nd)(*self.shape)
        self._dist_c = (ctypes.c_double * nd)(*self.distances)

    def _args(self):
        lib = self.lib
        return (_TORCH_REAL[self.dtype], len(self.shape), self._shape_c, self._dist_c, self.n_los, self.n,
                _ptr(lib, self.start), self.ss, _ptr(lib, self.end), self.es)

    def __call__(self, x, out=None):
        if tuple(x.shape) != self.shape or x.dtype != self.dtype:
            raise ValueError(f"LOS response is for shape {self.shape} {self.dtype}")
        out = torch.empty(self.n_los, dtype=self.dtype, device=x.device) if out is None else out
        lib = self.lib
        lib.check(lib.nb2_los_forward(*self._args(), _ptr(lib, x.contiguous()), _ptr(lib, out), _stream(x)), "nb2_los_forward")
        return out

    def adjoint(self, ybar, out=None):
        """``R^T ybar`` (the linear transpose the reference obtains with ``jax.linear_transpose``)."""
        out = torch.zeros(self.shape, dtype=self.dtype, device=ybar.device) if out is None else out.zero_()
        lib = self.lib
        lib.check(lib.nb2_los_adjoint(*self._args(), _ptr(lib, ybar.contiguous()), _ptr(lib, out), _stream(ybar)), "nb2_los_adjoint")
        return out


class LOSMetric:
    """``M t = A H f' R^T N^-1 R f' H A t + t`` on device tensors (xi block): the metric of the tomography
    configuration.  ``link_derivative``: ``f'(s)`` at the expansion point (position space; ``exp(s)`` for the
    demo's log-normal density), ``noise_cov_inv``: per-line inverse variance."""

    def __init__(self, cf, tables, los, link_derivative, noise_cov_inv, convention=None):
        self.cf, self.los, self.convention = cf, los, convention
        self.tab = cf.streaming_tables(tables)
        self.fp = link_derivative.to(device=cf.device, dtype=cf.dtype).contiguous()
        self.ninv = noise_cov_inv.to(device=cf.device, dtype=cf.dtype).contiguous()
        self._s = torch.empty(cf.shape, dtype=cf.dtype, device=cf.device)
        self._z = torch.empty(cf.shape, dtype=cf.dtype, device=cf.device)
        self._y = torch.empty(los.n_los, dtype=cf.dtype, device=cf.device)
        self._curv = torch.zeros((), dtype=torch.float64, device=cf.device)
        self.shape, self.dtype, self.device = tuple(cf.shape), cf.dtype, cf.device
        self.napply = 0

    def apply_with_curvature(self, t, out=None, pre=None, curv_out=None):
        if pre is not None:
            raise ValueError("LOSMetric does not carry the fused direction update")
        self.napply += 1
        cf = self.cf
        curv = self._curv if curv_out is None else curv_out
        s = cf.apply_xi(t, self.tab, out=self._s, convention=self.convention, offset=0.0, mul=self.fp)     # f' H A t
        y = self.los(s, out=self._y)                                                                     # R .
        y.mul_(self.ninv)                                                                                # N^-1 (n_los values)
        z = self.los.adjoint(y, out=self._z)                                                             # R^T .
        # f' rides on the first load of the adjoint transform (prologue field), A, + t and the dot on its last store
        epi = PointOp(self.tab.amp_mode, self.tab.amp, self.tab.azm * cf.dvol, add=t, dot=t, dot_out=curv,
                      amp_sym=getattr(self.tab, "sym", False))
        q = cf.plans[0].hartley(z, out=out, convention=self.convention, pro=PointOp(2, self.fp, 1.0), epi=epi)
        return q, curv

    def __call__(self, t, out=None):
        return self.apply_with_curvature(t, out=out)[0]
