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
    """O(N) direct evaluation of single output elements of the all-axes Hartley transform (SURVEY.md
    section 8d "sub-sampled checks"), for sizes where a full CPU transform is too slow or too large.
    ``idx_list``: iterable of index tuples.  The array is consumed slab by slab along its first axis
    (any float dtype; at most 16 MB of it is held in fp64 at a time), all requested elements at once:
    ``F[k] = sum_n0 e0[n0] (sum_rest a[n0, ...] e_rest[...])`` with fp64 phases ``exp(-2 pi i k n / N)``."""
    a = np.asarray(a)
    idx = [tuple(int(v) for v in i) for i in idx_list]
    ns = len(idx)
    shape = a.shape
    ph = [np.exp(-2j * np.pi * ((np.array([i[ax] for i in idx])[:, None] * np.arange(n)[None, :]) % n) / n)
          for ax, n in enumerate(shape)]                      # ph[ax][sample, n]
    acc = np.zeros(ns, dtype=np.complex128)
    if a.ndim == 1:
        acc = ph[0] @ a.astype(np.float64)
    else:
        rest = int(np.prod(shape[1:]))
        step = max(1, (2 << 20) // rest)
        for lo in range(0, shape[0], step):
            blk = a[lo:lo + step].astype(np.float64)           # (s, n1, ..., n_last)
            t = np.tensordot(blk, ph[-1], axes=([blk.ndim - 1], [1]))      # (s, n1, ..., sample)
            for ax in range(a.ndim - 2, 0, -1):
                # multiply by the phase of axis `ax` (per sample) and sum that axis away
                t = np.einsum("...ns,sn->...s", t, ph[ax])
            acc += np.einsum("ns,sn->s", t, ph[0][:, lo:lo + step])
    f = np.asarray(acc)
    return f.real + f.imag if convention == NON_CANONICAL else f.real - f.imag
