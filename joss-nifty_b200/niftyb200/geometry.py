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
        # O(n_bins) set-up tables, evaluated ONCE per grid on the host in NumPy like the reference does
        # (re/correlated_field.py:127-132): log_volume is a difference of nearly equal logarithms, which turns
        # a one-ulp difference of a device `log` into a 1e-9 relative change of the Wiener-process steps
        um = self.mode_lengths.cpu().numpy().copy()
        um[1:] = np.log(um[1:])
        um[1:] -= um[1]
        rel = torch.from_numpy(um).to(self.device)
        self.relative_log_mode_lengths = rel          # re/correlated_field.py:127-131
        self.log_volume = torch.from_numpy(um[2:] - um[1:-1]).to(self.device)         # :132

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
        self.nbins = int(sq.numel())
        self.mode_lengths = self._reference_mode_lengths(sq)

    def _reference_mode_lengths(self, unique_ksq):
        """Mode length of every bin with the reference's rounding: the reference evaluates
        ``sqrt(sum_j (min(k_j, N_j - k_j) * h_j)^2)`` in fp64 for every grid point and keeps, per bin, the
        largest of the (one-ulp apart) values (``rg_space.py:105-116``; ``re/correlated_field.py:55-72``:
        ``unique`` + the 1e-12 tolerance filter keep the last member of each cluster).  Same here, over the
        half grid ``0 .. N_j/2`` (the mirrored index takes no other values), slab by slab, with a scatter-max
        into the ksq table -- no N-sized array."""
        d = len(self.shape)
        h = self.hdist
        if d == 1:
            return torch.arange(self.shape[0] // 2 + 1, dtype=torch.float64, device=self.device) * h[0]
        if self.device.type != "cuda":
            return torch.from_numpy(self._reference_mode_lengths_host(unique_ksq.cpu().numpy())).to(self.device)
        # CUDA: fp64 sqrt / multiply / add are IEEE-exact there (torch's vectorised CPU sqrt is not)
        ax = [torch.arange(n // 2 + 1, dtype=torch.int64, device=self.device) for n in self.shape]
        sq = []
        for a, hh in zip(ax, h):
            v = a.to(torch.float64) * hh
            sq.append(v * v)                                                      # (k * h)^2 per axis
        table = torch.zeros(self.ksq_size, dtype=torch.float64, device=self.device)
        rest = int(np.prod([a.numel() for a in ax[1:]]))
        step = max(1, (1 << 22) // rest)
        for lo in range(0, ax[0].numel(), step):
            k0 = ax[0][lo:lo + step]
            acc_f, acc_i = sq[0][lo:lo + step], k0 * k0
            for a, f in zip(ax[1:], sq[1:]):                                      # ((f0 + f1) + f2): add.outer order
                acc_f = acc_f[..., None] + f
                acc_i = acc_i[..., None] + a * a
            table.scatter_reduce_(0, acc_i.reshape(-1), torch.sqrt(acc_f).reshape(-1), "amax", include_self=True)
        return table[unique_ksq]

    def _reference_mode_lengths_host(self, unique_ksq):
        """NumPy twin of :meth:`_reference_mode_lengths` (CPU tensors: the emulator tests)."""
        ax = [np.arange(n // 2 + 1, dtype=np.int64) for n in self.shape]
        sq = []
        for a, hh in zip(ax, self.hdist):
            v = a.astype(np.float64) * hh
            sq.append(v * v)
        table = np.zeros(self.ksq_size, dtype=np.float64)
        rest = int(np.prod([a.size for a in ax[1:]]))
        step = max(1, (1 << 22) // rest)
        for lo in range(0, ax[0].size, step):
            k0 = ax[0][lo:lo + step]
            acc_f, acc_i = sq[0][lo:lo + step], k0 * k0
            for a, f in zip(ax[1:], sq[1:]):
                acc_f = np.add.outer(acc_f, f)
                acc_i = np.add.outer(acc_i, a * a)
            idx, val = acc_i.ravel(), np.sqrt(acc_f).ravel()
            order = np.argsort(idx, kind="stable")
            idx, val = idx[order], val[order]
            first = np.flatnonzero(np.r_[True, idx[1:] != idx[:-1]])
            np.maximum.at(table, idx[first], np.maximum.reduceat(val, first))
        return table[unique_ksq]

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
