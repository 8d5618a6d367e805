"""Oracle: the sampling line-of-sight response (test infrastructure).

Restates ``_los`` / ``SamplingCartesianGridLOS`` (``/root/reference/src/re/extra/sampling_los.py:15-91``) in NumPy,
including the algorithm of ``jax.scipy.ndimage.map_coordinates(order=1, mode='constant', cval=nan)`` it relies on
(jax is absent here): per axis ``lower = floor(c)``, weights ``(1 - (c - lower), c - lower)`` on ``lower`` and
``lower + 1``; BOTH corners are always visited and a corner outside the grid contributes ``cval`` times its weight
(so a sample on the last grid plane yields NaN: ``nan * 0``).  The adjoint is the transpose of that linear map
(``jax.linear_transpose`` in the reference); corners outside the grid have no adjoint.
"""

import itertools

import numpy as np


def _points(start, end, shape, distances, n_sampling_points):
    shape = np.asarray(shape, dtype=np.float64)
    l2i = ((shape - 1) / shape) / np.asarray(distances, dtype=np.float64)          # :18
    start_iloc, end_iloc = start * l2i, end * l2i                                   # :19-20
    ddi = (end_iloc - start_iloc) / n_sampling_points                               # :21
    adi = np.arange(0, n_sampling_points) + 0.5                                     # :22
    dist = np.linalg.norm(end - start)                                              # :23
    pp = start_iloc[:, np.newaxis] + ddi[:, np.newaxis] * adi[np.newaxis]           # :24
    return pp, dist


def _corners(x_shape, pp):
    lower = np.floor(pp)
    wu = pp - lower
    lower = lower.astype(np.int64)
    for corner in itertools.product((0, 1), repeat=len(x_shape)):
        idx = [lower[i] + c for i, c in enumerate(corner)]
        w = np.ones(pp.shape[1])
        inside = np.ones(pp.shape[1], dtype=bool)
        for i, c in enumerate(corner):
            w = w * (wu[i] if c else 1.0 - wu[i])
            inside &= (idx[i] >= 0) & (idx[i] < x_shape[i])
        yield tuple(np.where(inside, ix, 0) for ix in idx), w, inside


def los(x, start, end, *, distances, shape, n_sampling_points):
    """One line (``_los``, :15-28)."""
    pp, dist = _points(np.asarray(start, float), np.asarray(end, float), shape, distances, n_sampling_points)
    val = np.zeros(pp.shape[1])
    for idx, w, inside in _corners(x.shape, pp):
        val = val + w * np.where(inside, x[idx], np.nan)
    return val.sum() * (dist / n_sampling_points)


class SamplingCartesianGridLOS:
    """``SamplingCartesianGridLOS`` (:31-91), ``interpolation_order=1``."""

    def __init__(self, start, end, *, shape, distances, n_sampling_points=500):
        self.start, self.end = np.atleast_2d(np.asarray(start, float)), np.atleast_2d(np.asarray(end, float))
        self.shape, self.n = tuple(shape), int(n_sampling_points)
        self.distances = np.broadcast_to(np.asarray(distances, float), (len(self.shape),))
        self.n_los = max(self.start.shape[0], self.end.shape[0])

    def _se(self, l):
        return self.start[l if self.start.shape[0] > 1 else 0], self.end[l if self.end.shape[0] > 1 else 0]

    def __call__(self, x):
        return np.array([los(x, *self._se(l), distances=self.distances, shape=self.shape, n_sampling_points=self.n)
                         for l in range(self.n_los)])

    def adjoint(self, ybar):
        out = np.zeros(self.shape)
        for l in range(self.n_los):
            pp, dist = _points(*self._se(l), self.shape, self.distances, self.n)
            for idx, w, inside in _corners(self.shape, pp):
                np.add.at(out, tuple(ix[inside] for ix in idx), w[inside] * ybar[l] * (dist / self.n))
        return out
