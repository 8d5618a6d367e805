"""Oracle: harmonic transforms at the ``ducc_dispatch`` seam (test infrastructure).

Restates ``/root/reference/src/ducc_dispatch.py`` (the SciPy fallback branch,
``:38-58,88-92``) and the scaling rules of ``HartleyOperator._apply_cartesian``
(``src/operators/harmonic_operators.py:208-217``).
"""

import os

import numpy as np
import scipy.fft

NON_CANONICAL = "non_canonical_hartley"
CANONICAL = "canonical_hartley"

_nthreads = 1


def set_nthreads(n):
    """``ducc_dispatch.set_nthreads`` (``src/ducc_dispatch.py:33-35``)."""
    global _nthreads
    _nthreads = int(n)


def nthreads():
    return _nthreads


def _workers():
    return _nthreads if _nthreads > 0 else os.cpu_count()


def fftn(a, axes=None):
    """``_scipy_fftn`` (``src/ducc_dispatch.py:38-39``): unnormalised forward c2c."""
    return scipy.fft.fftn(a, axes=axes, workers=_workers())


def ifftn(a, axes=None):
    """``_scipy_ifftn`` (``src/ducc_dispatch.py:42-43``): backward c2c scaled 1/N."""
    return scipy.fft.ifftn(a, axes=axes, workers=_workers())


def hartley(a, axes=None, convention=NON_CANONICAL):
    """``_scipy_hartley`` (``src/ducc_dispatch.py:46-50``).

    Genuine (non-separable) multi-D Hartley transform, unnormalised:
    ``F = fftn(a, axes)``; non-canonical -> ``F.real + F.imag``,
    canonical -> ``F.real - F.imag``.
    """
    tmp = scipy.fft.fftn(a, axes=axes, workers=_workers())
    if convention == NON_CANONICAL:
        return tmp.real + tmp.imag
    if convention == CANONICAL:
        return tmp.real - tmp.imag
    raise ValueError(f"invalid hartley convention {convention!r}")


def vdot(a, b):
    """``_scipy_vdot`` (``src/ducc_dispatch.py:53-58``): sum(conj(a)*b)."""
    return np.vdot(a, b)


def hartley_operator(x, axes, fct, convention=NON_CANONICAL):
    """``HartleyOperator.apply`` (``harmonic_operators.py:200-217``).

    Complex input: real and imaginary parts are transformed separately
    (``:202-204``).  ``fct`` is the scalar volume factor; ``fct == 1`` skips the
    multiply (``:217``).
    """
    x = np.asarray(x)
    if np.iscomplexobj(x):
        return (hartley_operator(x.real, axes, fct, convention)
                + 1j * hartley_operator(x.imag, axes, fct, convention))
    tmp = hartley(x, axes=axes, convention=convention)
    return tmp if fct == 1 else tmp * fct


def direct_hartley_samples(a, idx_list, convention=NON_CANONICAL):
    """O(N) direct evaluation of single output elements of the all-axes Hartley
    transform (SURVEY.md section 8d "sub-sampled checks"): used at sizes where a full
    CPU transform is too slow/large.  ``idx_list``: iterable of index tuples."""
    a = np.asarray(a, dtype=np.float64)
    out = []
    grids = [np.arange(n, dtype=np.float64) for n in a.shape]
    for idx in idx_list:
        # separable phases, accumulated in extended fashion: cos(sum) -/+ sin(sum)
        ph = [2 * np.pi * ((int(k) * g) % n) / n for k, g, n in zip(idx, grids, a.shape)]
        e = [np.exp(-1j * p) for p in ph]
        acc = a.astype(np.complex128)
        for ax in range(a.ndim - 1, -1, -1):
            acc = np.tensordot(acc, e[ax], axes=([ax], [0]))
        f = complex(acc)
        out.append(f.real + f.imag if convention == NON_CANONICAL else f.real - f.imag)
    return np.array(out)
