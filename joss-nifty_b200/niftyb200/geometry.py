"""Regular-grid harmonic geometry: mode lengths, power bins and the tables the kernels use.

Follows ``RGSpace`` / ``PowerSpace`` of the numpy stack
(``/root/reference/src/domains/rg_space.py:77-150``, ``src/domains/power_space.py:170-198``)
and ``get_fourier_mode_distributor`` / ``_make_grid`` of the JAX stack
(``src/re/correlated_field.py:30-142``).

B200 layout decision: the reference stores an int64 ``pindex`` per grid point (8.6 GB at
1024^3).  For isotropic grids (equal harmonic pixel size on every axis) the bin of a mode is
a function of the integer ``ksq = sum_j min(k_j, N_j - k_j)^2`` alone, so the kernels index a
dense ``ksq`` table (<= 0.8 M entries at 1024^3, L2 resident) instead.  Anisotropic grids fall
back to an explicit (lazily built) index.
"""

import numpy as np
import torch


def _as_tuple(x):
    return (int(x),) if np.isscalar(x) else tuple(int(v) for v in x)


class RGGrid:
    """Harmonic partner geometry of a regular position-space grid."""

    def __init__(self, shape, distances, device="cpu"):
        self.shape = _as_tuple(shape)
        d = len(self.shape)
        self.distances = tuple(float(v) for v in np.broadcast_to(np.asarray(distances, dtype=np.float64), (d,)))
        self.device = torch.device(device)
        self.total_volume = float(np.prod(np.array(self.shape, dtype=np.float64) * np.array(self.distances)))
        # harmonic pixel sizes 1/(n_i d_i)  (rg_space.py:77-78)
        self.hdist = tuple(1.0 / (n * dd) for n, dd in zip(self.shape, self.distances))
        self.isotropic = all(h == self.hdist[0] for h in self.hdist)
        self._pindex = None
        if self.isotropic:
            self._build_isotropic()
        else:
            self._build_general()
        um = self.mode_lengths
        rel = um.clone()
        rel[1:] = torch.log(rel[1:])
        rel[1:] -= rel[1].clone()
        self.relative_log_mode_lengths = rel          # re/correlated_field.py:127-131
        self.log_volume = rel[2:] - rel[1:-1]         # :132

    # -- isotropic: everything from integer ksq, no N-sized array -------------------
    def _axis_sq(self, n):
        k = torch.arange(n // 2 + 1, dtype=torch.int64, device=self.device)
        sq = k * k
        mult = torch.full_like(k, 2)
        mult[0] = 1
        if n % 2 == 0:
            mult[-1] = 1
        if n == 1:
            mult[0] = 1
        return sq, mult

    def _build_isotropic(self):
        sq, mult = self._axis_sq(self.shape[0])
        mult = mult.to(torch.float64)
        for n in self.shape[1:]:
            s2, m2 = self._axis_sq(n)
            sq = (sq[:, None] + s2[None, :]).reshape(-1)
            mult = (mult[:, None] * m2.to(torch.float64)[None, :]).reshape(-1)
            # compress after every axis to keep the temporaries small
            size = int(sq.max().item()) + 1
            cnt = torch.bincount(sq, weights=mult, minlength=size)
            sq = torch.nonzero(cnt, as_tuple=False).reshape(-1)
            mult = cnt[sq]
        if len(self.shape) == 1:
            size = int(sq.max().item()) + 1
        self.ksq_size = int(sum((n // 2) ** 2 for n in self.shape)) + 1
        self.unique_ksq = sq                                   # sorted int64
        self.mode_multiplicity = mult.round().to(torch.int64)  # points per bin
        self.mode_lengths = torch.sqrt(sq.to(torch.float64)) * self.hdist[0]
        self.nbins = int(sq.numel())

    def ksq_table(self, amp, dtype):
        """Dense table T[ksq] = amp[bin(ksq)] for the kernels' amp_mode 1."""
        if not self.isotropic:
            raise ValueError("ksq tables exist for isotropic grids only")
        t = torch.zeros(self.ksq_size, dtype=dtype, device=amp.device)
        t[self.unique_ksq.to(amp.device)] = amp.to(dtype)
        return t

    # -- general: explicit index (reference algorithm) ------------------------------------
    def _k_length_array(self):
        """``RGSpace._get_dist_array`` (rg_space.py:105-116)."""
        res = None
        for n, h in zip(self.shape, self.hdist):
            k = torch.arange(n, dtype=torch.float64, device=self.device)
            k = torch.minimum(k, n - k) * h
            res = k * k if res is None else (res[..., None] + (k * k))
        return torch.sqrt(res)

    def _build_general(self):
        kl = self._k_length_array()
        um = torch.unique(kl)
        tol = 1e-12 * um[-1]
        keep = torch.diff(torch.cat((um, 2 * um[-1:]))) > tol    # re/correlated_field.py:69-72
        um = um[keep]
        bounds = 0.5 * (um[:-1] + um[1:])
        idx = torch.searchsorted(bounds, kl.reshape(-1).contiguous()).reshape(kl.shape)
        cnt = torch.bincount(idx.reshape(-1), minlength=um.numel())
        if (cnt == 0).any():
            raise RuntimeError("invalid harmonic mode(s) encountered")
        self._pindex = idx
        self.mode_lengths = um
        self.mode_multiplicity = cnt
        self.nbins = int(um.numel())
        self.unique_ksq = None
        self.ksq_size = 0

    @property
    def pindex(self):
        """Bin index per grid point (``PowerSpace.pindex``, power_space.py:178); built lazily."""
        if self._pindex is None:
            ksq = None
            for n in self.shape:
                k = torch.arange(n, dtype=torch.int64, device=self.device)
                k = torch.minimum(k, n - k)
                ksq = k * k if ksq is None else (ksq[..., None] + k * k)
            lut = torch.full((self.ksq_size,), -1, dtype=torch.int64, device=self.device)
            lut[self.unique_ksq] = torch.arange(self.nbins, device=self.device)
            self._pindex = lut[ksq]
        return self._pindex

    def expand(self, amp, dtype):
        """``amp[pindex]`` (PowerDistributor.times, distributors.py:114-119) as a full field."""
        return amp.to(dtype)[self.pindex.to(amp.device)]
