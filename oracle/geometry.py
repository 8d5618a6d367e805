"""Oracle: regular-grid harmonic geometry and the power (mode-length) index.

Test infrastructure.  Restates
* ``RGSpace._get_dist_array`` / ``get_unique_k_lengths``
  (``/root/reference/src/domains/rg_space.py:105-150``),
* ``PowerSpace.__init__`` natural binning (``src/domains/power_space.py:170-198``),
* ``get_fourier_mode_distributor`` and ``_make_grid``
  (``src/re/correlated_field.py:30-142``).
"""

from collections import namedtuple

import numpy as np


def _as_tuple(shape):
    return (int(shape),) if np.isscalar(shape) else tuple(int(s) for s in shape)


def harmonic_distances(shape, distances):
    """Harmonic-space pixel distances ``1/(n_i * d_i)`` (``rg_space.py:77-78``,
    ``re/correlated_field.py:55``)."""
    shape = _as_tuple(shape)
    distances = np.broadcast_to(np.asarray(distances, dtype=np.float64), (len(shape),))
    return 1.0 / (np.array(shape, dtype=np.float64) * distances)


def k_length_array(shape, distances):
    """``RGSpace._get_dist_array`` on the harmonic partner (``rg_space.py:105-116``)
    == the ``m_length`` block of ``get_fourier_mode_distributor``
    (``re/correlated_field.py:55-65``)."""
    shape = _as_tuple(shape)
    hd = harmonic_distances(shape, distances)
    res = np.arange(shape[0], dtype=np.float64)
    res = np.minimum(res, shape[0] - res) * hd[0]
    if len(shape) == 1:
        return res
    res *= res
    for i in range(1, len(shape)):
        tmp = np.arange(shape[i], dtype=np.float64)
        tmp = np.minimum(tmp, shape[i] - tmp) * hd[i]
        tmp *= tmp
        res = np.add.outer(res, tmp)
    return np.sqrt(res)


def unique_k_lengths_np(shape, distances):
    """``RGSpace.get_unique_k_lengths`` (``rg_space.py:123-150``), numpy stack."""
    shape = _as_tuple(shape)
    hd = harmonic_distances(shape, distances)
    d = len(shape)
    if d == 1:
        return np.arange(shape[0] // 2 + 1, dtype=np.float64) * hd[0]
    if np.all(hd == hd[0]):
        maxdist = np.asarray(shape) // 2
        tmp = np.zeros(int(np.sum(maxdist * maxdist)) + 1, dtype=bool)
        t2 = np.arange(maxdist[0] + 1, dtype=np.int64)
        t2 *= t2
        for i in range(1, d):
            t3 = np.arange(maxdist[i] + 1, dtype=np.int64)
            t3 *= t3
            t2 = np.add.outer(t2, t3)
        tmp[t2] = True
        return np.sqrt(np.nonzero(tmp)[0]) * hd[0]
    tmp = np.unique(k_length_array(shape, distances))
    tol = 1e-12 * tmp[-1]
    return tmp[np.diff(np.r_[tmp, 2 * tmp[-1]]) > tol]


PowerIndex = namedtuple("PowerIndex", "pindex k_lengths rho unique_k")


def power_index_np(shape, distances):
    """``PowerSpace.__init__`` with natural binning (``power_space.py:170-198``)."""
    kl = k_length_array(shape, distances)
    uk = unique_k_lengths_np(shape, distances)
    tbb = 0.5 * (uk[:-1] + uk[1:])
    pindex = np.searchsorted(tbb, kl)
    nbin = len(tbb) + 1
    rho = np.bincount(pindex.ravel(), minlength=nbin)
    if (rho == 0).any():
        raise ValueError("empty bins detected")
    k_lengths = np.bincount(pindex.ravel(), weights=kl.ravel(), minlength=nbin).astype(np.float64) / rho
    return PowerIndex(pindex, k_lengths, rho, uk)


def fourier_mode_distributor_isotropic(shape, distances):
    """Same result as :func:`fourier_mode_distributor` for grids with equal harmonic pixel sizes, from the
    integer squared mode lengths ``ksq = sum_j min(k_j, N_j - k_j)^2`` (the device the reference's own numpy
    stack uses in ``RGSpace.get_unique_k_lengths``, ``rg_space.py:130-141``) instead of a sort of N floats.
    Only used above ``_ISOTROPIC_SHORTCUT_SIZE`` points, where the float sort takes minutes; equality with the
    restated reference algorithm is asserted in tests/test_oracle_golden.py."""
    shape = _as_tuple(shape)
    hd = harmonic_distances(shape, distances)
    ksq = None
    for n in shape:
        k = np.arange(n, dtype=np.int64)
        k = np.minimum(k, n - k) ** 2
        ksq = k if ksq is None else np.add.outer(ksq, k)
    m_count_all = np.bincount(ksq.ravel())
    uniq = np.nonzero(m_count_all)[0]
    lut = np.zeros(m_count_all.size, dtype=np.int64)
    lut[uniq] = np.arange(uniq.size)
    # mode length of a bin with the reference's rounding: the largest of the fp64 values
    # sqrt(((k0 h)^2 + (k1 h)^2) + ...) over the bin's points (``unique`` + the tolerance filter of
    # ``:69-72`` keep the last member of every cluster of nearly equal values); plane by plane
    um = np.zeros(uniq.size, dtype=np.float64)
    f = []
    for n, h in zip(shape, hd):
        v = np.arange(n, dtype=np.float64)
        v = np.minimum(v, n - v) * h
        f.append(v * v)
    idx = lut[ksq]
    for i0 in range(shape[0] // 2 + 1):              # planes i0 and n0 - i0 hold the same values
        acc = f[0][i0]
        for fj in f[1:]:
            acc = np.add.outer(acc, fj)
        np.maximum.at(um, idx[i0].ravel(), np.sqrt(acc).ravel())
    return idx, um, m_count_all[uniq]


_ISOTROPIC_SHORTCUT_SIZE = 1 << 24


def fourier_mode_distributor(shape, distances):
    """``get_fourier_mode_distributor`` (``re/correlated_field.py:30-81``)."""
    hd = harmonic_distances(shape, distances)
    if np.prod(_as_tuple(shape)) > _ISOTROPIC_SHORTCUT_SIZE and len(hd) > 1 and np.all(hd == hd[0]):
        return fourier_mode_distributor_isotropic(shape, distances)
    m_length = k_length_array(shape, distances)
    um = np.unique(m_length)
    tol = 1e-12 * um[-1]
    um = um[np.diff(np.append(um, 2 * um[-1])) > tol]
    binbounds = 0.5 * (um[:-1] + um[1:])
    m_length_idx = np.searchsorted(binbounds, m_length)
    m_count = np.bincount(m_length_idx.ravel(), minlength=um.size)
    if np.any(m_count == 0) or um.shape != m_count.shape:
        raise RuntimeError("invalid harmonic mode(s) encountered")
    return m_length_idx, um, m_count


Grid = namedtuple("Grid", "shape total_volume distances power_distributor mode_multiplicity "
                          "mode_lengths relative_log_mode_lengths log_volume")


_grid_cache = {}


def make_grid(shape, distances):
    """``_make_grid`` (``re/correlated_field.py:108-142``), fourier harmonic type (memoised: the grid
    depends on the shape and the distances only)."""
    shape = _as_tuple(shape)
    distances = tuple(float(d) for d in np.broadcast_to(distances, (len(shape),)).astype(np.float64))
    key = (shape, distances)
    if key not in _grid_cache:
        if len(_grid_cache) > 8:
            _grid_cache.clear()
        _grid_cache[key] = _make_grid(shape, distances)
    return _grid_cache[key]


def _make_grid(shape, distances):
    totvol = float(np.prod(np.array(shape) * np.array(distances)))
    idx, m_length, m_count = fourier_mode_distributor(shape, distances)
    um = m_length.copy()
    um[1:] = np.log(um[1:])
    um[1:] -= um[1]
    assert um[0] == 0.0
    log_vol = um[2:] - um[1:-1]
    return Grid(shape, totvol, distances, idx, m_count, m_length, um, log_vol)
