"""FFTOperator / HartleyOperator / HarmonicTransformOperator on the GPU backend.

Mirror of ``/root/reference/src/operators/harmonic_operators.py:34-123,126-217,326-380`` for regular
grids: same modes (TIMES, ADJOINT_TIMES, INVERSE_TIMES, ADJOINT_INVERSE_TIMES), same volume
factors (``:213-217``: domain ``scalar_dvol`` for TIMES|ADJOINT_TIMES, target ``scalar_dvol``
for the inverse modes), complex input transformed part by part (``:202-206``).  Fields are
plain arrays (numpy in -> numpy out, torch device tensors in -> device tensors out); the
``Field``/``DomainTuple`` containers of the reference stay in the reference.
"""

import numpy as np
import torch

from .transforms import PointOp, _CPLX_OF, _TORCH_CPLX, _TORCH_REAL, get_plan, device as _device


class RGSpace:
    """Minimal ``RGSpace`` (``src/domains/rg_space.py:25-210``): shape, distances, harmonic flag."""

    def __init__(self, shape, distances=None, harmonic=False):
        self.shape = (int(shape),) if np.isscalar(shape) else tuple(int(s) for s in shape)
        if distances is None:
            distances = 1.0 if harmonic else tuple(1.0 / s for s in self.shape)   # rg_space.py:58-64
        self.distances = tuple(float(d) for d in np.broadcast_to(np.asarray(distances, float), (len(self.shape),)))
        self.harmonic = bool(harmonic)

    @property
    def scalar_dvol(self):
        """``rg_space.py:77-82``."""
        return float(np.prod(self.distances))

    def get_default_codomain(self):
        """``rg_space.py:152-166``: distances ``1/(n d)``, harmonic flag flipped."""
        return RGSpace(self.shape, tuple(1.0 / (n * d) for n, d in zip(self.shape, self.distances)),
                       not self.harmonic)

    def check_codomain(self, codomain):
        """``rg_space.py:168-197``."""
        if not isinstance(codomain, RGSpace):
            raise TypeError("domain is not a RGSpace")
        if self.shape != codomain.shape:
            raise AttributeError("The shapes of domain and codomain must be identical.")
        if self.harmonic == codomain.harmonic:
            raise AttributeError("domain.harmonic and codomain.harmonic must not be the same.")
        if not np.allclose(np.array(self.distances) * np.array(codomain.distances) * np.array(self.shape), 1.0,
                           rtol=1e-7, atol=0):
            raise AttributeError("The grid-distances of domain and codomain do not match.")


class HartleyOperator:
    """``HartleyOperator`` (``harmonic_operators.py:126-217``) for a single RGSpace, optionally
    embedded at ``axes`` of a larger array of ``full_shape`` (other axes are batch)."""

    TIMES, ADJOINT_TIMES, INVERSE_TIMES, ADJOINT_INVERSE_TIMES = 1, 2, 4, 8

    def __init__(self, domain, target=None, full_shape=None, axes=None):
        if not isinstance(domain, RGSpace):
            raise TypeError("HartleyOperator only works on RGSpaces")
        self.domain = domain
        self.target = domain.get_default_codomain() if target is None else target
        domain.check_codomain(self.target)
        self.target.check_codomain(domain)
        self.full_shape = domain.shape if full_shape is None else tuple(full_shape)
        self.axes = tuple(range(len(self.full_shape))) if axes is None else tuple(axes)
        if tuple(self.full_shape[a] for a in self.axes) != domain.shape:
            raise ValueError("axes/full_shape do not match the domain")

    def _fct(self, mode):
        if mode & (self.TIMES | self.ADJOINT_TIMES):
            return self.domain.scalar_dvol
        if mode & (self.INVERSE_TIMES | self.ADJOINT_INVERSE_TIMES):
            return self.target.scalar_dvol
        raise ValueError("invalid operator mode specified")

    def _real(self, x, fct):
        plan = get_plan(x.shape, self.axes, x.dtype)
        return plan.hartley(x, epi=None if fct == 1 else PointOp(scale=fct))

    def apply(self, x, mode):
        if mode not in (1, 2, 4, 8):
            raise ValueError("invalid operator mode specified")
        host = not isinstance(x, torch.Tensor)
        t = torch.as_tensor(np.ascontiguousarray(x)) if host else x
        if tuple(t.shape) != self.full_shape:
            raise ValueError("The operator's and field's domains don't match.")
        if not (t.is_complex() or t.dtype in (torch.float32, torch.float64)):
            t = t.to(torch.float64)
        t = t.to(_device()).contiguous()
        fct = self._fct(mode)
        if t.is_complex():
            y = torch.complex(self._real(t.real.contiguous(), fct), self._real(t.imag.contiguous(), fct))
        else:
            y = self._real(t, fct)
        return y.cpu().numpy() if host else y

    def times(self, x):
        return self.apply(x, self.TIMES)

    def adjoint_times(self, x):
        return self.apply(x, self.ADJOINT_TIMES)

    def inverse_times(self, x):
        return self.apply(x, self.INVERSE_TIMES)

    def adjoint_inverse_times(self, x):
        return self.apply(x, self.ADJOINT_INVERSE_TIMES)

    __call__ = times


class FFTOperator(HartleyOperator):
    """``FFTOperator`` (``harmonic_operators.py:34-123``): full complex FFT between a position-space and a
    harmonic RGSpace; the result is always complex (``:52-56``).  A field on a position space is transformed
    with ``fftn`` (factor 1), a field on a harmonic space with ``ifftn`` times the number of cells (``:108-114``);
    the volume factor is the domain's ``scalar_dvol`` for TIMES | ADJOINT_TIMES and the target's for the
    inverse modes (``:119-122``).  All four modes (``_all_ops``, ``:60``)."""

    def __init__(self, domain, target=None, full_shape=None, axes=None):
        if not isinstance(domain, RGSpace):
            raise TypeError("FFTOperator only works on RGSpaces")
        super().__init__(domain, target, full_shape, axes)

    def apply(self, x, mode):
        if mode not in (1, 2, 4, 8):
            raise ValueError("invalid operator mode specified")
        host = not isinstance(x, torch.Tensor)
        t = torch.as_tensor(np.ascontiguousarray(x)) if host else x
        if tuple(t.shape) != self.full_shape:
            raise ValueError("The operator's and field's domains don't match.")
        if not t.is_complex():
            t = t.to(torch.complex64 if t.dtype == torch.float32 else torch.complex128)
        t = t.to(_device()).contiguous()
        # the space the INPUT lives on: the domain for TIMES / ADJOINT_INVERSE_TIMES, else the target
        src = self.domain if mode & (self.TIMES | self.ADJOINT_INVERSE_TIMES) else self.target
        ncells = int(np.prod(src.shape))
        fct = self._fct(mode)
        plan = get_plan(t.shape, self.axes, t.dtype)
        if src.harmonic:          # harmonic -> position: ifftn * ncells == unnormalised backward transform
            y = plan.c2c(t, forward=False, scale=fct)
        else:
            y = plan.c2c(t, forward=True, scale=fct)
        del ncells
        return y.cpu().numpy() if host else y


class HarmonicTransformOperator(HartleyOperator):
    """``HarmonicTransformOperator`` (``harmonic_operators.py:326-380``): harmonic -> position,
    TIMES | ADJOINT_TIMES only (``:375``); on RG grids it is the Hartley operator (``:369-370``)."""

    def __init__(self, domain, target=None, full_shape=None, axes=None):
        if not domain.harmonic:
            raise TypeError("HarmonicTransformOperator only works on a harmonic space")
        super().__init__(domain, target, full_shape, axes)

    def apply(self, x, mode):
        if mode not in (self.TIMES, self.ADJOINT_TIMES):
            raise NotImplementedError("HarmonicTransformOperator supports TIMES and ADJOINT_TIMES only")
        return super().apply(x, mode)
