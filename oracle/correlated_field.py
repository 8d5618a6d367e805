"""Oracle: correlated-field model, its hot tail (forward apply) and VJP (adjoint).

Test infrastructure.  NumPy restatement of ``CorrelatedFieldMaker`` of the JAX
stack (``/root/reference/src/re/correlated_field.py:313-704``; jax absent here),
pinned against the importable numpy stack
(``src/library/correlated_fields.py:764-823`` ``finalize``) via ``tests/golden``.

The hot tail is ``correlated_field`` (``re/correlated_field.py:690-693``)::

    ea   = outer_i( normed_amp_i(p)[power_distributor_i] )      # :670-688
    cf_h = azm(p) * ea * p[prefix + "xi"]                        # :692
    y    = offset_mean + prod_i harmonic_dvol_i * hartley_i(cf_h) # :663-668,693

with ``normed_amp = amp.at[1:].mul(1/azm)`` (``:605-611``) and ``amp[0] = totvol``
(``:303``).  The adjoint is the VJP of that function (``jax.vjp`` in the
reference, ``re/likelihood.py:299,573-583``); numpy-stack twin:
``PowerDistributor._adjoint_times`` = bincount (``src/operators/distributors.py:105-112``).
"""

import numpy as np

from . import amplitude as _amp
from . import geometry as _geo
from . import transforms as _tr


class _Fluct:
    def __init__(self, grid, kind, priors, prefix, np_kind):
        self.grid, self.kind, self.priors, self.prefix, self.np_kind = grid, kind, priors, prefix, np_kind

    def domain(self):
        d = {}
        if self.kind == "nonparametric":
            d[self.prefix + "fluctuations"] = ()
            d[self.prefix + "loglogavgslope"] = ()
            if self.priors["flexibility"] is not None and self.grid.log_volume.size > 0:
                d[self.prefix + "flexibility"] = ()
                if self.priors["asperity"] is not None:
                    d[self.prefix + "asperity"] = ()
                d[self.prefix + "spectrum"] = (self.grid.log_volume.size, 2)
        else:
            for k in ("scale", "cutoff", "loglogslope"):
                d[self.prefix + k] = ()
        return d

    def amplitude(self, p):
        pr, g, pf = self.priors, self.grid, self.prefix
        if self.kind == "nonparametric":
            flu = pr["fluctuations"](p[pf + "fluctuations"])
            slp = pr["loglogavgslope"](p[pf + "loglogavgslope"])
            flx = asp = sxi = None
            if pr["flexibility"] is not None and g.log_volume.size > 0:
                flx = pr["flexibility"](p[pf + "flexibility"])
                if pr["asperity"] is not None:
                    asp = pr["asperity"](p[pf + "asperity"])
                sxi = p[pf + "spectrum"]
            return _amp.non_parametric_amplitude(g, flu, slp, flx, asp, sxi, kind=self.np_kind)
        return _amp.matern_amplitude(
            g, pr["scale"](p[pf + "scale"]), pr["cutoff"](p[pf + "cutoff"]),
            pr["loglogslope"](p[pf + "loglogslope"]), pr["renormalize_amplitude"], kind=self.np_kind)


class CorrelatedFieldMaker:
    """``re.correlated_field.CorrelatedFieldMaker`` (``:313-704``), numpy restatement."""

    def __init__(self, prefix):
        self._prefix = prefix
        self._azm = None
        self._offset_mean = None
        self._fl = []

    def set_amplitude_total_offset(self, offset_mean, offset_std):
        """``:557-582``; ``offset_std`` = (mean, std) of a log-normal prior."""
        self._offset_mean = offset_mean
        self._azm = _amp.lognormal_prior(*offset_std)

    def add_fluctuations(self, shape, distances, fluctuations, loglogavgslope, flexibility=None,
                         asperity=None, prefix="", harmonic_type="fourier",
                         non_parametric_kind="amplitude"):
        """``:343-447``."""
        grid = _geo.make_grid(shape, distances)
        pri = dict(
            fluctuations=_amp.lognormal_prior(*fluctuations),
            loglogavgslope=_amp.normal_prior(*loglogavgslope),
            flexibility=None if flexibility is None else _amp.lognormal_prior(*flexibility),
            asperity=None if asperity is None else _amp.lognormal_prior(*asperity),
        )
        self._fl.append(_Fluct(grid, "nonparametric", pri, self._prefix + prefix, non_parametric_kind))

    def add_fluctuations_matern(self, shape, distances, scale, cutoff, loglogslope,
                                renormalize_amplitude, prefix="", harmonic_type="fourier",
                                non_parametric_kind="amplitude"):
        """``:449-555``."""
        grid = _geo.make_grid(shape, distances)
        pri = dict(scale=_amp.lognormal_prior(*scale), cutoff=_amp.lognormal_prior(*cutoff),
                   loglogslope=_amp.normal_prior(*loglogslope),
                   renormalize_amplitude=renormalize_amplitude)
        self._fl.append(_Fluct(grid, "matern", pri, self._prefix + prefix, non_parametric_kind))

    def finalize(self):
        return CorrelatedField(self)


class CorrelatedField:
    """Result of ``finalize`` (``re/correlated_field.py:640-704``)."""

    def __init__(self, mk):
        self.prefix = mk._prefix
        self.offset_mean = mk._offset_mean
        self._azm = mk._azm
        self.fl = list(mk._fl)
        self.grids = [f.grid for f in self.fl]
        self.shape = tuple(s for g in self.grids for s in g.shape)
        self.domain = {self.prefix + "xi": self.shape, self.prefix + "zeromode": ()}
        for f in self.fl:
            self.domain.update(f.domain())
        # axes of each sub-grid inside the full excitation shape (:646-650)
        self.axes, n = [], 0
        for g in self.grids:
            self.axes.append(tuple(range(n, n + len(g.shape))))
            n += len(g.shape)

    def azm(self, p):
        return self._azm(p[self.prefix + "zeromode"])

    def normalized_amplitudes(self, p):
        """``get_normalized_amplitudes`` (``:597-611``): ``amp.at[1:].mul(1/azm)``."""
        azm = self.azm(p)
        out = []
        for f in self.fl:
            a = f.amplitude(p).copy()
            a[1:] *= 1.0 / azm
            out.append(a)
        return out

    def outer_amplitude(self, amps):
        """``outer_amplitude`` (``:681-688``): gather + tensordot(axes=0)."""
        outer = amps[0][self.grids[0].power_distributor]
        for a, g in zip(amps[1:], self.grids[1:]):
            outer = np.tensordot(outer, a[g.power_distributor], axes=0)
        return outer

    def harmonic_transform(self, x, convention=_tr.NON_CANONICAL):
        """``outer_harmonic_transform`` (``:663-668``): one Hartley per sub-grid, chained."""
        out = x
        for g, ax in zip(self.grids, self.axes):
            out = (1.0 / g.total_volume) * _tr.hartley(out, axes=ax, convention=convention)
        return out

    def __call__(self, p, convention=_tr.NON_CANONICAL):
        """``correlated_field`` (``:690-693``)."""
        ea = self.outer_amplitude(self.normalized_amplitudes(p))
        cf_h = self.azm(p) * ea * p[self.prefix + "xi"]
        return self.offset_mean + self.harmonic_transform(cf_h, convention)

    # ---- the hot tail with the amplitude tables given (what the CUDA path accelerates) ----
    def apply(self, amps, azm, xi, convention=_tr.NON_CANONICAL, offset=None):
        """Forward apply for fixed tables: ``offset + HT(azm * outer(amps) * xi)``."""
        off = self.offset_mean if offset is None else offset
        return off + self.harmonic_transform(azm * self.outer_amplitude(amps) * xi, convention)

    def adjoint(self, amps, azm, ybar, xi=None, convention=_tr.NON_CANONICAL):
        """VJP of :meth:`apply` w.r.t. ``xi`` and, if ``xi`` is given, the tables.

        Hartley is symmetric, so its transpose is the same chain in reverse order
        (``harmonic_operators.py:213-217``: TIMES|ADJOINT_TIMES share one branch).
        Returns ``(xibar, [ampbar_g], azmbar, offsetbar)``; table cotangents are the
        binned reductions of ``distributors.py:105-112``.
        """
        h = ybar
        for g, ax in reversed(list(zip(self.grids, self.axes))):
            h = (1.0 / g.total_volume) * _tr.hartley(h, axes=ax, convention=convention)
        ea = self.outer_amplitude(amps)
        xibar = azm * ea * h
        if xi is None:
            return xibar, None, None, float(np.sum(ybar))
        u = xi * h                       # d y / d (azm*ea) contracted with ybar
        azmbar = float(np.sum(ea * u))
        ampbars = []
        for i, (g, ax) in enumerate(zip(self.grids, self.axes)):
            w = azm * u
            # multiply by the other sub-grids' expanded amplitudes (broadcast on their axes)
            for j, (a, gg) in enumerate(zip(amps, self.grids)):
                if j == i:
                    continue
                shp = [1] * len(self.shape)
                for k, axk in enumerate(self.axes[j]):
                    shp[axk] = gg.shape[k]
                w = w * a[gg.power_distributor].reshape(shp)
            red_axes = tuple(k for k in range(len(self.shape)) if k not in ax)
            w = w.sum(axis=red_axes) if red_axes else w
            ampbars.append(np.bincount(g.power_distributor.ravel(), weights=w.ravel(),
                                       minlength=amps[i].size))
        return xibar, ampbars, azmbar, float(np.sum(ybar))
