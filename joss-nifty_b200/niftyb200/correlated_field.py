"""Correlated-field model with the hot tail on the GPU.

Interface follows ``nifty8.re.correlated_field.CorrelatedFieldMaker``
(``/root/reference/src/re/correlated_field.py:311-704``) -- ``add_fluctuations``,
``add_fluctuations_matern``, ``set_amplitude_total_offset``, ``finalize`` -- and
``SimpleCorrelatedField`` of the numpy stack
(``src/library/correlated_fields_simple.py:38-170``).  Parameters are a dict
``{prefix+"xi", prefix+"zeromode", prefix+"fluctuations", ...}`` of standard-normal
excitations exactly as in the reference (``finalize`` ``:640-704``).

What runs where:
* O(n_bins) amplitude chain (integrated Wiener process, normalisation; ``:277-304``,
  ``src/re/gauss_markov.py:105-121``): fp64 torch ops on the device (cold path,
  SURVEY.md section 8f N1),
* hot tail ``offset + (1/V) hartley(azm * amp[pindex] * xi)`` (``:690-693``) and its adjoint:
  one fused 3-pass CUDA transform (:mod:`niftyb200.transforms`), the gather riding on the
  first load, scale/offset on the last store.
"""

import math

import numpy as np
import torch

from . import _lib, config
from .geometry import RGGrid
from .transforms import Plan, PointOp, device as _device


# ---------------------------------------------------------------- priors (re/num/stats_distributions.py:41-88)
def lognormal_moments(mean, std):
    if mean <= 0.0:
        raise ValueError(f"`mean` must be greater zero; got {mean!r}")
    if std <= 0.0:
        raise ValueError(f"`std` must be greater zero; got {std!r}")
    logstd = math.sqrt(math.log1p((std / mean) ** 2))
    return math.log(mean) - 0.5 * logstd ** 2, logstd


def _host_float(v):
    """Scalar leaf (NumPy / Python number or 0-d tensor on any device) as a Python float."""
    if isinstance(v, torch.Tensor):
        return float(v.detach().cpu())
    return float(np.asarray(v))


class _ScalarPrior:
    """Standard-normal excitation -> parameter (re/num/stats_distributions.py:41-88), callable on torch
    tensors like the plain lambdas, plus value and derivative as host floats for the scalar leaves (the
    amplitude-chain kernels take the transformed scalars by value)."""

    def __init__(self, kind, a, b):
        self.kind, self.a, self.b = kind, float(a), float(b)      # normal: a + b xi; lognormal: exp(a + b xi)

    def __call__(self, xi):
        if self.kind == "normal":
            return self.a + self.b * xi
        return torch.exp(self.a + self.b * xi) if isinstance(xi, torch.Tensor) else math.exp(self.a + self.b * xi)

    def value(self, xi):
        return float(self(float(xi)))

    def grad(self, xi):
        """d value / d xi"""
        return self.b if self.kind == "normal" else self.b * self.value(xi)


def normal_prior(mean, std):
    return _ScalarPrior("normal", mean, std)


def lognormal_prior(mean, std):
    lm, ls = lognormal_moments(mean, std)
    return _ScalarPrior("lognormal", lm, ls)


def _prior(spec, kind):
    if spec is None or callable(spec):
        return spec
    if len(spec) != 2:
        raise TypeError(f"prior specification must be (mean, std); got {spec!r}")
    return (lognormal_prior if kind == "lognormal" else normal_prior)(*spec)


# ---------------------------------------------------------------- amplitude chain (torch, fp64)
def _integrated_wiener_process(xi, sigma, dt, asperity):
    """``integrated_wiener_process`` (src/re/gauss_markov.py:105-121) with x0 = 0;
    returns the integrated component only."""
    amp = sigma * torch.sqrt(dt)
    dev = amp * xi[:, 1]                                       # increments of the rate
    lvl = amp * xi[:, 0] * torch.sqrt(dt * dt / 12.0 + (0.0 if asperity is None else asperity))
    lvl = lvl + 0.5 * dt * dev
    zero = torch.zeros(1, dtype=xi.dtype, device=xi.device)
    rate = torch.cumsum(torch.cat((zero, dev)), 0)
    lvl = torch.cat((zero, lvl + dt * rate[:-1]))
    return torch.cumsum(lvl, 0)


class _Fluctuations:
    def __init__(self, grid, kind, priors, prefix, np_kind):
        self.grid, self.kind, self.priors, self.prefix, self.np_kind = grid, kind, priors, prefix, np_kind.lower()
        if self.np_kind not in ("amplitude", "power"):
            raise ValueError(f"invalid kind specified {np_kind!r}")

    def domain(self):
        d, pf, g = {}, self.prefix, self.grid
        if self.kind == "nonparametric":
            d[pf + "fluctuations"] = ()
            d[pf + "loglogavgslope"] = ()
            if self.priors["flexibility"] is not None and g.log_volume.numel() > 0:
                d[pf + "flexibility"] = ()
                if self.priors["asperity"] is not None:
                    d[pf + "asperity"] = ()
                d[pf + "spectrum"] = (int(g.log_volume.numel()), 2)
        else:
            for k in ("scale", "cutoff", "loglogslope"):
                d[pf + k] = ()
        return d

    @property
    def on_kernel(self):
        """The chain runs in the CUDA kernels of csrc/nb2_amplitude.cuh: non-parametric spectrum with the
        standard (mean, std) priors.  Matern spectra and user-supplied prior callables use the torch ops below."""
        if self.kind != "nonparametric":
            return False
        keys = ["fluctuations", "loglogavgslope"]
        if self.priors["flexibility"] is not None and self.grid.log_volume.numel() > 0:
            keys.append("flexibility")
            if self.priors["asperity"] is not None:
                keys.append("asperity")
        return all(isinstance(self.priors[k], _ScalarPrior) for k in keys)

    def _chain_inputs(self, p, azm):
        g, pr, pf = self.grid, self.priors, self.prefix
        lat = {k: _host_float(p[pf + k]) for k in ("fluctuations", "loglogavgslope")}
        flex = pr["flexibility"] is not None and g.log_volume.numel() > 0
        asp = flex and pr["asperity"] is not None
        if flex:
            lat["flexibility"] = _host_float(p[pf + "flexibility"])
        if asp:
            lat["asperity"] = _host_float(p[pf + "asperity"])
        xi = None
        if flex:
            v = p[pf + "spectrum"]
            xi = (v if isinstance(v, torch.Tensor) else torch.as_tensor(np.asarray(v))).to(device=g.device, dtype=torch.float64)
            xi = xi.contiguous()
        sc = dict(flu=pr["fluctuations"].value(lat["fluctuations"]), slope=pr["loglogavgslope"].value(lat["loglogavgslope"]),
                  flx=pr["flexibility"].value(lat["flexibility"]) if flex else -1.0,
                  asp=pr["asperity"].value(lat["asperity"]) if asp else -1.0, azm=float(azm))
        return lat, xi, sc

    def _geom_args(self):
        g = self.grid
        if not hasattr(g, "_mult_f64"):
            g._mult_f64 = g.mode_multiplicity.to(torch.float64).contiguous()
        return g.relative_log_mode_lengths.contiguous(), g.log_volume.contiguous(), g._mult_f64

    def amplitude_kernel(self, p, azm, table=None, table_idx=None):
        """NORMALISED amplitude per power bin (amp[0] = totvol, amp[1:] / azm; re/correlated_field.py:277-304,
        605-611) by the device kernel; returns (amp, context for :meth:`vjp_kernel`)."""
        lib = _lib.get()
        g = self.grid
        rel, dt, mult = self._geom_args()
        lat, xi, sc = self._chain_inputs(p, azm)
        nbins = int(rel.numel())
        scratch = torch.empty(2 * nbins + 2, dtype=torch.float64, device=g.device)
        amp = torch.empty(nbins, dtype=torch.float64, device=g.device)
        kind = 1 if self.np_kind == "power" else 0
        from .transforms import _ptr, _stream
        rc = lib.nb2_amplitude_forward(nbins, _ptr(lib, rel), _ptr(lib, dt) if nbins > 2 else None, _ptr(lib, mult),
                                       float(g.total_volume), _ptr(lib, xi), sc["flu"], sc["slope"], sc["flx"], sc["asp"],
                                       sc["azm"], kind, _ptr(lib, scratch), _ptr(lib, amp),
                                       _ptr(lib, table), _ptr(lib, table_idx),
                                       1 if (table is not None and table.dtype == torch.float32) else 0, _stream(amp))
        lib.check(rc, "nb2_amplitude_forward")
        return amp, dict(lat=lat, xi=xi, sc=sc, scratch=scratch, amp=amp, kind=kind)

    def vjp_kernel(self, ctx, ampbar):
        """Cotangents of this sub-space's latents for the cotangent ``ampbar`` of its normalised amplitudes:
        ``({leaf: cotangent}, azmbar through the chain)``."""
        lib = _lib.get()
        g, pr, pf = self.grid, self.priors, self.prefix
        rel, dt, mult = self._geom_args()
        nbins = int(rel.numel())
        xi, sc, lat = ctx["xi"], ctx["sc"], ctx["lat"]
        xibar = torch.zeros_like(xi) if xi is not None else None
        sbar = torch.zeros(5, dtype=torch.float64, device=g.device)
        work = torch.empty(2 * nbins, dtype=torch.float64, device=g.device)
        from .transforms import _ptr, _stream
        ab = ampbar.to(device=g.device, dtype=torch.float64).contiguous()
        rc = lib.nb2_amplitude_vjp(nbins, _ptr(lib, rel), _ptr(lib, dt) if nbins > 2 else None, _ptr(lib, mult),
                                   float(g.total_volume), _ptr(lib, xi), sc["flu"], sc["slope"], sc["flx"], sc["asp"],
                                   sc["azm"], ctx["kind"], _ptr(lib, ctx["scratch"]), _ptr(lib, ctx["amp"]), _ptr(lib, ab),
                                   _ptr(lib, work), _ptr(lib, xibar), _ptr(lib, sbar), _stream(ab))
        lib.check(rc, "nb2_amplitude_vjp")
        sb = sbar.cpu().numpy()
        out = {pf + "fluctuations": sb[0] * pr["fluctuations"].grad(lat["fluctuations"]),
               pf + "loglogavgslope": sb[1] * pr["loglogavgslope"].grad(lat["loglogavgslope"])}
        if "flexibility" in lat:
            out[pf + "flexibility"] = sb[2] * pr["flexibility"].grad(lat["flexibility"])
            out[pf + "spectrum"] = xibar
        if "asperity" in lat:
            out[pf + "asperity"] = sb[3] * pr["asperity"].grad(lat["asperity"])
        return out, float(sb[4])

    def jvp_kernel(self, ctx, t, dazm):
        """Tangent of this sub-space's normalised amplitudes for the latent tangents ``t`` (dict of the
        sub-space's leaves; missing = 0) and the tangent ``dazm`` of the zero-mode amplitude."""
        import ctypes
        lib = _lib.get()
        g, pr, pf = self.grid, self.priors, self.prefix
        rel, dt, mult = self._geom_args()
        nbins = int(rel.numel())
        xi, sc, lat = ctx["xi"], ctx["sc"], ctx["lat"]

        def ds(name):
            v = t.get(pf + name)
            return 0.0 if (v is None or name not in lat) else _host_float(v) * pr[name].grad(lat[name])
        dsc = (ctypes.c_double * 5)(ds("fluctuations"), ds("loglogavgslope"), ds("flexibility"), ds("asperity"), float(dazm))
        dxi = t.get(pf + "spectrum") if xi is not None else None
        if dxi is not None:
            dxi = (dxi if isinstance(dxi, torch.Tensor) else torch.as_tensor(np.asarray(dxi))).to(device=g.device,
                                                                                                  dtype=torch.float64).contiguous()
        work = torch.empty(2 * nbins, dtype=torch.float64, device=g.device)
        damp = torch.empty(nbins, dtype=torch.float64, device=g.device)
        from .transforms import _ptr, _stream
        rc = lib.nb2_amplitude_jvp(nbins, _ptr(lib, rel), _ptr(lib, dt) if nbins > 2 else None, _ptr(lib, mult),
                                   float(g.total_volume), _ptr(lib, xi), sc["flu"], sc["slope"], sc["flx"], sc["asp"],
                                   sc["azm"], ctx["kind"], _ptr(lib, ctx["scratch"]), _ptr(lib, ctx["amp"]), _ptr(lib, dxi),
                                   dsc, _ptr(lib, work), _ptr(lib, damp), _stream(damp))
        lib.check(rc, "nb2_amplitude_jvp")
        return damp

    # ---- chain derivatives for what the kernels do not cover (Matern spectra, user prior callables): the torch
    # ops of :meth:`amplitude` differentiated by autograd -- O(n_bins), once per nonlinear evaluation
    def leaf_names(self):
        return [k for k in self.domain()]

    def torch_context(self, pd, azm):
        """Context of :meth:`vjp_torch` / :meth:`jvp_torch`: this sub-space's leaves as fp64 device tensors."""
        return dict(torch=True, leaves={k: pd[k].detach().to(torch.float64) for k in self.leaf_names()}, azm=float(azm))

    def _normalised(self, leaves, azm):
        a = self.amplitude(leaves)
        return torch.cat((a[:1], a[1:] / azm))

    def vjp_torch(self, ctx, ampbar):
        names = list(ctx["leaves"])
        leaves = {k: ctx["leaves"][k].clone().requires_grad_(True) for k in names}
        azm = torch.tensor(ctx["azm"], dtype=torch.float64, device=self.grid.device, requires_grad=True)
        an = self._normalised(leaves, azm)
        ab = ampbar.to(device=an.device, dtype=torch.float64)
        grads = torch.autograd.grad(an, [leaves[k] for k in names] + [azm], grad_outputs=ab, allow_unused=True)
        out = {}
        for k, gk in zip(names, grads[:-1]):
            gk = torch.zeros_like(leaves[k]) if gk is None else gk
            out[k] = gk if gk.dim() else float(gk)
        return out, 0.0 if grads[-1] is None else float(grads[-1])

    def jvp_torch(self, ctx, t, dazm):
        names = list(ctx["leaves"])
        prim = tuple(ctx["leaves"][k] for k in names) + (torch.tensor(ctx["azm"], dtype=torch.float64,
                                                                       device=self.grid.device),)
        tang = []
        for k in names:
            v = t.get(k)
            if v is None:
                tang.append(torch.zeros_like(ctx["leaves"][k]))
            else:
                v = v if isinstance(v, torch.Tensor) else torch.as_tensor(np.asarray(v))
                tang.append(v.to(device=self.grid.device, dtype=torch.float64).reshape(ctx["leaves"][k].shape))
        tang.append(torch.tensor(float(dazm), dtype=torch.float64, device=self.grid.device))

        def fn(*args):
            return self._normalised(dict(zip(names, args[:-1])), args[-1])
        return torch.autograd.functional.jvp(fn, prim, tuple(tang))[1]

    def vjp(self, ctx, ampbar):
        return self.vjp_torch(ctx, ampbar) if ctx.get("torch") else self.vjp_kernel(ctx, ampbar)

    def jvp(self, ctx, t, dazm):
        return self.jvp_torch(ctx, t, dazm) if ctx.get("torch") else self.jvp_kernel(ctx, t, dazm)

    def amplitude(self, p):
        """Un-normalised amplitude per power bin (fp64 device tensor)."""
        g, pr, pf = self.grid, self.priors, self.prefix
        totvol = g.total_volume
        mult = g.mode_multiplicity.to(torch.float64)
        if self.kind == "nonparametric":
            # non_parametric_amplitude.correlate (re/correlated_field.py:277-304)
            flu = pr["fluctuations"](p[pf + "fluctuations"])
            slope = pr["loglogavgslope"](p[pf + "loglogavgslope"])
            rel = g.relative_log_mode_lengths
            ln_spec = slope * rel
            if pr["flexibility"] is not None and g.log_volume.numel() > 0:
                flx = pr["flexibility"](p[pf + "flexibility"])
                asp = pr["asperity"](p[pf + "asperity"]) if pr["asperity"] is not None else None
                tw = _integrated_wiener_process(p[pf + "spectrum"], flx, g.log_volume, asp)
                tw = torch.cat((torch.zeros(1, dtype=tw.dtype, device=tw.device), tw))
                ln_spec = ln_spec + (tw - tw[-1] * (rel / rel[-1]))       # _remove_slope :144-147
            spec = torch.exp(ln_spec)
            if self.np_kind == "amplitude":
                norm = torch.sqrt(torch.sum(mult[1:] * spec[1:] ** 2)) / math.sqrt(totvol)
                amp = flu * (math.sqrt(totvol) / norm) * spec
            else:
                norm = torch.sqrt(torch.sum(mult[1:] * spec[1:])) / math.sqrt(totvol)
                amp = flu * (math.sqrt(totvol) / norm) * torch.sqrt(spec)
        else:
            # matern_amplitude.correlate (re/correlated_field.py:186-213)
            scale = pr["scale"](p[pf + "scale"])
            cutoff = pr["cutoff"](p[pf + "cutoff"])
            slope = pr["loglogslope"](p[pf + "loglogslope"])
            spec = torch.exp(0.25 * slope * torch.log1p((g.mode_lengths / cutoff) ** 2))
            norm = 1.0
            if pr["renormalize_amplitude"]:
                pw = 4 if self.np_kind == "amplitude" else 2
                norm = torch.sqrt(torch.sum(mult[1:] * spec[1:] ** pw)) / math.sqrt(totvol)
            amp = scale * (math.sqrt(totvol) / norm) * spec
            if self.np_kind == "power":
                amp = torch.cat((amp[:1], torch.sqrt(amp[1:])))
        amp = amp.clone()
        amp[0] = totvol if not (self.kind == "matern" and self.np_kind == "power") else math.sqrt(totvol)
        return amp


class CorrelatedFieldMaker:
    """``nifty8.re.CorrelatedFieldMaker`` (re/correlated_field.py:311-704) for the GPU backend."""

    def __init__(self, prefix, dtype=torch.float64):
        self._prefix = prefix
        self._azm = None
        self._offset_mean = None
        self._fluctuations = []
        self._dtype = dtype

    def set_amplitude_total_offset(self, offset_mean, offset_std):
        """``:548-575``; ``offset_std`` = (mean, std) of a log-normal prior or a callable."""
        zm = offset_std
        if not callable(zm):
            if zm is None or len(zm) != 2:
                raise TypeError(f"`offset_std` of invalid type {type(zm)!r}")
            zm = lognormal_prior(*zm)
        self._offset_mean = float(offset_mean)
        self._azm = zm

    def add_fluctuations(self, shape, distances, fluctuations, loglogavgslope, flexibility=None,
                         asperity=None, prefix="", harmonic_type="fourier", non_parametric_kind="amplitude"):
        """``:343-450``."""
        if harmonic_type.lower() != "fourier":
            raise NotImplementedError("only harmonic_type='fourier' (regular grids) is on the GPU path")
        grid = RGGrid(shape, distances, _device())
        pri = dict(fluctuations=_prior(fluctuations, "lognormal"), loglogavgslope=_prior(loglogavgslope, "normal"),
                   flexibility=_prior(flexibility, "lognormal"), asperity=_prior(asperity, "lognormal"))
        self._fluctuations.append(_Fluctuations(grid, "nonparametric", pri, self._prefix + prefix,
                                                non_parametric_kind))

    def add_fluctuations_matern(self, shape, distances, scale, cutoff, loglogslope, renormalize_amplitude,
                                prefix="", harmonic_type="fourier", non_parametric_kind="amplitude"):
        """``:452-546``."""
        if harmonic_type.lower() != "fourier":
            raise NotImplementedError("only harmonic_type='fourier' (regular grids) is on the GPU path")
        grid = RGGrid(shape, distances, _device())
        pri = dict(scale=_prior(scale, "lognormal"), cutoff=_prior(cutoff, "lognormal"),
                   loglogslope=_prior(loglogslope, "normal"), renormalize_amplitude=bool(renormalize_amplitude))
        self._fluctuations.append(_Fluctuations(grid, "matern", pri, self._prefix + prefix, non_parametric_kind))

    @property
    def azm(self):
        if self._azm is None:
            raise NotImplementedError("You need to set the `amplitude_total_offset` first")
        return self._azm

    def finalize(self):
        """``:640-704``."""
        if not self._fluctuations:
            raise NotImplementedError("add fluctuations first")
        _ = self.azm
        return CorrelatedField(self._prefix, self._offset_mean, self._azm, list(self._fluctuations), self._dtype)


class Tables:
    """Everything the hot tail needs for one expansion point: device-resident."""

    def __init__(self, amp_mode, amp, azm, amps, sym=False):
        self.amp_mode, self.amp, self.azm, self.amps = amp_mode, amp, float(azm), amps
        # amp_mode-2 fields built from per-axis mode lengths satisfy field[k] == field[-k] along every
        # axis (rg_space.py:105-116): the kernels may read each mirror pair of rows once (NB2_AMP_SYMMETRIC)
        self.sym = bool(sym) and amp_mode == 2


class CorrelatedField:
    """Finalised model: ``cf(p)`` == ``correlated_field(p)`` (re/correlated_field.py:690-693)."""

    def __init__(self, prefix, offset_mean, azm, fluctuations, dtype):
        self.prefix, self.offset_mean, self._azm, self.fl, self.dtype = prefix, offset_mean, azm, fluctuations, dtype
        self.grids = [f.grid for f in self.fl]
        self.shape = tuple(s for g in self.grids for s in g.shape)
        self.domain = {prefix + "xi": self.shape, prefix + "zeromode": ()}
        for f in self.fl:
            self.domain.update(f.domain())
        self.axes, n = [], 0
        for g in self.grids:
            self.axes.append(tuple(range(n, n + len(g.shape))))
            n += len(g.shape)
        self.device = _device()
        self.plans = [Plan(self.shape, ax, dtype) for ax in self.axes]
        self.single = len(self.grids) == 1
        # (space) x (1-D space), all powers of two: ONE joint fast-path plan whose row pass combines the
        # two rows of a mirror pair into the product of the two Hartley transforms
        # (NB2_HARTLEY_PRODUCT_LAST) -- 3 passes instead of one transform per sub-space
        self.joint = None
        if len(self.grids) == 2 and len(self.grids[1].shape) == 1 and len(self.shape) <= 3:
            jp = Plan(self.shape, None, dtype)
            if jp.is_fast:
                self.joint = jp
        # joint harmonic volume factor prod_g 1/V_g (re/correlated_field.py:654,663-668)
        self.dvol = float(np.prod([1.0 / g.total_volume for g in self.grids]))

    # ---- parameters ---------------------------------------------------------------
    def _as_dev(self, v, dtype=torch.float64):
        if isinstance(v, torch.Tensor):
            return v.to(device=self.device, dtype=dtype)
        return torch.as_tensor(np.asarray(v), dtype=dtype).to(self.device)

    def init(self, seed=0, dtype=None):
        """Standard-normal excitations for every parameter (``Model.init``)."""
        gen = np.random.default_rng(seed)
        return {k: gen.standard_normal(shp) for k, shp in self.domain.items()}

    def azm(self, p):
        return self._azm(self._as_dev(p[self.prefix + "zeromode"]))

    def normalized_amplitudes(self, p):
        """``get_normalized_amplitudes`` (``:598-611``): ``amp.at[1:].mul(1/azm)``, fp64, device."""
        amps, azm, _ = self._amplitudes(p)
        return amps, azm

    def _amplitudes(self, p):
        """(normalised amplitudes, azm, per-sub-space chain context or None).  Sub-spaces with a non-parametric
        spectrum and standard priors run the chain in ONE kernel (csrc/nb2_amplitude.cuh);
        Matern spectra / user prior callables use torch ops."""
        zl = p[self.prefix + "zeromode"]
        if isinstance(self._azm, _ScalarPrior):
            azm = self._azm.value(_host_float(zl))
        else:
            azm = self._azm(self._as_dev(zl))
        out, ctxs, pd = [], [], None
        for f in self.fl:
            if f.on_kernel and isinstance(azm, float):
                a, ctx = f.amplitude_kernel(p, azm)
                out.append(a)
                ctxs.append(ctx)
                continue
            if pd is None:
                pd = {k: self._as_dev(v) for k, v in p.items() if not k.endswith("xi")}
            a = f.amplitude(pd)
            out.append(torch.cat((a[:1], a[1:] / azm)))
            ctxs.append(f.torch_context(pd, azm) if isinstance(azm, float) else None)
        return out, azm, ctxs

    def tables(self, p):
        """Device tables for the expansion point ``p`` (cold path, once per nonlinear step)."""
        amps, azm, ctxs = self._amplitudes(p)
        tab = self.tables_from_amplitudes(amps, float(azm))
        tab.chain = ctxs          # what vjp_latent needs to continue through the amplitude chain
        return tab

    def jvp_latent(self, t, p, convention=None, tab=None, mul=None, out=None):
        """Directional derivative of the model along the latent tangent ``t`` (dict over the leaves; missing = 0):
        ``d cf(p)[t]`` -- ``lin.jac(t)`` / ``jax.jvp`` (golden ``jvp`` vectors).  With
        ``A = azm prod_g amp_g[pindex_g]``:  ``dy = (1/V) H(dA xi + A dxi) = (1/V) H(A (R xi + dxi))``,
        ``R = dA / A = dazm/azm + sum_g damp_g/amp_g`` (amp_g[0] is a constant), so the tangent costs ONE fused
        transform after an O(N) combine.  ``mul``: fused multiplier on the last store (likelihood weight)."""
        tab = self.tables(p) if tab is None else tab
        if getattr(tab, "chain", None) is None or any(c is None for c in tab.chain) or not isinstance(self._azm, _ScalarPrior):
            raise NotImplementedError("jvp_latent: the zero-mode prior must be a standard (mean, std) prior")
        zl = _host_float(p[self.prefix + "zeromode"])
        azm = self._azm.value(zl)
        tz = t.get(self.prefix + "zeromode")
        dazm = 0.0 if tz is None else _host_float(tz) * self._azm.grad(zl)
        ratio = None
        iso = self.single and self.grids[0].isotropic and self.plans[0].is_fast
        for f, g, ctx, a in zip(self.fl, self.grids, tab.chain, tab.amps):
            damp = f.jvp(ctx, t, dazm)
            r = damp / a
            r[0] = 0.0                                           # amp[0] = totvol: constant
            if iso:     # the gather through the ksq table, as for the amplitude itself: no N-sized index
                ratio = self.plans[0].expand_table(g.ksq_table(r + dazm / azm, self.dtype))
                break
            e = g.expand(r, torch.float64)
            if ratio is None:
                ratio = e
            else:
                shp = list(ratio.shape) + [1] * e.dim()
                ratio = ratio.reshape(shp) + e
        if not iso:
            ratio = (ratio + dazm / azm).to(self.dtype)
        xi = self._as_dev(p[self.prefix + "xi"], self.dtype)
        v = ratio * xi
        dxi = t.get(self.prefix + "xi")
        if dxi is not None:
            v = v + self._as_dev(dxi, self.dtype)
        return self.apply_xi(v.contiguous(), tab, out=out, convention=convention, offset=0.0, mul=mul)

    def vjp_latent(self, ybar, p, convention=None, tab=None):
        """Cotangent of EVERY latent leaf of the model for the position-space cotangent ``ybar`` -- the
        Jacobian adjoint the reference gets from ``jax.vjp`` / ``Linearization.jac.adjoint_times``
        (re/likelihood.py:299,573-583; golden ``adj_*`` vectors): the hot tail's VJP (one transform with the
        binned reduction, :meth:`vjp`), then the amplitude chain's VJP kernel per sub-space and the priors'
        derivatives.  Returns ``{leaf: cotangent}`` (device tensors for ``xi`` / ``spectrum``, floats for the
        scalar leaves) and the cotangent of ``offset_mean`` under the key ``"offset_mean"``."""
        tab = self.tables(p) if tab is None else tab
        if getattr(tab, "chain", None) is None or any(c is None for c in tab.chain) or not isinstance(self._azm, _ScalarPrior):
            raise NotImplementedError("vjp_latent: the zero-mode prior must be a standard (mean, std) prior")
        xi = self._as_dev(p[self.prefix + "xi"], self.dtype).contiguous()
        yb = self._as_dev(ybar, self.dtype).contiguous()
        xibar, ampbars, azmbar, offbar = self.vjp(yb, tab, xi, convention=convention)
        out = {self.prefix + "xi": xibar, "offset_mean": offbar}
        for f, ctx, ab in zip(self.fl, tab.chain, ampbars):
            leaves, azm_chain = f.vjp(ctx, ab)
            out.update(leaves)
            azmbar += azm_chain
        zl = _host_float(p[self.prefix + "zeromode"])
        out[self.prefix + "zeromode"] = azmbar * self._azm.grad(zl)
        return out

    def tables_from_amplitudes(self, amps, azm):
        amps = [self._as_dev(a) for a in amps]
        if self.single and self.grids[0].isotropic and self.plans[0].is_fast:
            return Tables(1, self.grids[0].ksq_table(amps[0], self.dtype), azm, amps)
        full = None
        for a, g in zip(amps, self.grids):
            e = g.expand(a, torch.float64)
            full = e if full is None else torch.tensordot(full, e, dims=0)    # outer_amplitude :681-688
        return Tables(2, full.to(self.dtype).contiguous(), azm, amps, sym=True)

    def streaming_tables(self, tab):
        """Tables with the gather resolved once: ``field = azm * amp[pindex]`` as a full-size
        array (one extra N*s of HBM, streamed by the passes instead of a per-element L2 gather).
        Pays off whenever the same expansion point is applied many times (every CG solve)."""
        if tab.amp_mode != 1:
            return tab
        field = self.plans[0].expand_table(tab.amp, scale=tab.azm)
        st = Tables(2, field, 1.0, tab.amps, sym=True)
        st.azm_value = tab.azm        # the factor folded into the field
        return st

    # ---- hot tail ---------------------------------------------------------------------
    @property
    def fuses_direction_update(self):
        """True when the first pass of :meth:`apply_xi` is a strided pass of the fast path, which can
        carry the CG direction update ``xi <- pre[0] + pre[1] * xi`` (``nb2_pointop.pre_add``)."""
        pl = self.plans[0] if self.joint is None else self.joint
        nax = sum(1 for a in pl.axes if self.shape[a] > 1)
        return (pl.kind == 1 and nax >= 2) or (pl.kind == 2 and nax >= 1)

    def apply_xi(self, xi, tab, out=None, convention=None, offset=None, mul=None, pre=None):
        """``offset + prod_g (1/V_g) hartley_g(azm * amp[pindex] * xi)`` on device tensors.

        ``pre = (r, beta)``: first ``xi <- r + beta * xi`` IN PLACE (the conjugate-gradient direction
        update, conjugate_gradient.py:124), fused into the first pass; needs
        :attr:`fuses_direction_update` and ``out is not xi``."""
        off = self.offset_mean if offset is None else offset
        pro = PointOp(tab.amp_mode, tab.amp, tab.azm, amp_sym=getattr(tab, "sym", False))
        if pre is not None:
            if not self.fuses_direction_update:
                raise ValueError("this field's transform cannot carry the fused direction update")
            pro.pre_add, pro.pre_out = pre[0], xi
            if isinstance(pre[1], torch.Tensor):       # beta lives on the device (device-resident CG)
                pro.pre_scale_dev = pre[1]
            else:
                pro.pre_scale = float(pre[1])
        if self.single:
            return self.plans[0].hartley(xi, out=out, convention=convention, pro=pro,
                                         epi=PointOp(scale=self.dvol, add_scalar=off, mul=mul))
        if self.joint is not None:
            return self.joint.hartley(xi, out=out, convention=convention, pro=pro, product_last=True,
                                      epi=PointOp(scale=self.dvol, add_scalar=off, mul=mul))
        cur = self.plans[0].hartley(xi, out=out, convention=convention, pro=pro)
        for pl in self.plans[1:-1]:
            cur = pl.hartley(cur, out=cur, convention=convention)
        return self.plans[-1].hartley(cur, out=cur, convention=convention,
                                      epi=PointOp(scale=self.dvol, add_scalar=off, mul=mul))

    def adjoint_xi(self, ybar, tab, out=None, convention=None, add=None, dot=None, dot_out=None):
        """xi-block of the VJP: ``azm * amp[pindex] * prod_g (1/V_g) hartley_g(ybar)`` (Hartley is
        symmetric: harmonic_operators.py:213-217).  Optional fused ``+ add`` and ``dot``."""
        epi = PointOp(tab.amp_mode, tab.amp, tab.azm * self.dvol, add=add, dot=dot, dot_out=dot_out,
                      amp_sym=getattr(tab, "sym", False))
        if self.single:
            return self.plans[0].hartley(ybar, out=out, convention=convention, epi=epi)
        if self.joint is not None:     # the sub-space transforms act on different axes: they commute
            return self.joint.hartley(ybar, out=out, convention=convention, epi=epi, product_last=True)
        cur = self.plans[-1].hartley(ybar, out=out, convention=convention)
        for pl in reversed(self.plans[1:-1]):
            cur = pl.hartley(cur, out=cur, convention=convention)
        return self.plans[0].hartley(cur, out=cur, convention=convention, epi=epi)

    def vjp(self, ybar, tab, xi, convention=None):
        """Full cotangent of the hot tail for ``ybar`` (position space) at excitations ``xi``:
        ``(xibar, [ampbar_g], azmbar, offsetbar)`` -- the xi block as in :meth:`adjoint_xi` plus the
        table cotangents, i.e. the binned reductions of ``PowerDistributor.adjoint_times``
        (``distributors.py:105-112``) that feed the VJP of the O(n_bins) amplitude chain.

        Single isotropic sub-space on the fast path: ONE transform whose epilogue also accumulates
        ``xi * h`` into the ksq bins (one extra read of ``xi``).  Anything else: the same algebra with
        torch reductions over the explicit ``pindex`` (cold path)."""
        amps = tab.amps
        if self.single and self.grids[0].isotropic and self.plans[0].is_fast:
            g = self.grids[0]
            # Order-independent accumulation (two-limb fixed point, integer atomics: nb2_pointop.bin_fixed) --
            # the cotangent of the spectrum is bit-reproducible from run to run.  Scale from the rigorous bound
            # |H(y)_i| <= sqrt(N) ||y||_2 on every transform value.
            nbins = self.plans[0].ksq_size
            scale = float(tab.azm) * self.dvol
            bound = 2.0 * (float(torch.linalg.vector_norm(xi, ord=float("inf"))) * math.sqrt(ybar.numel())
                           * float(torch.linalg.vector_norm(ybar)) * abs(scale))
            fscale = 2.0 ** (42 - math.ceil(math.log2(bound))) if 0.0 < bound < float("inf") else 1.0
            fixed = torch.zeros(2 + 2 * nbins, dtype=torch.int64, device=self.device)
            epi = PointOp(tab.amp_mode, tab.amp, scale, bin_x=xi, bin_fixed=fixed, bin_fixed_scale=fscale,
                          amp_sym=getattr(tab, "sym", False))
            xibar = self.plans[0].hartley(ybar, convention=convention, epi=epi)
            limbs = fixed[2:].view(nbins, 2).to(torch.float64)
            bins = (limbs[:, 0] + limbs[:, 1] * 2.0 ** -40) / fscale
            if int(fixed[0]) != 0 or not bound < float("inf"):     # non-finite input
                bins = torch.full_like(bins, float("nan"))
            # bins[ksq] = sum_i xi_i * (azm' * h_i) over the points of that ksq, azm' = tab.azm
            per_bin = bins[g.unique_ksq.to(self.device)]
            true_azm = float(tab.azm) if tab.amp_mode == 1 else float(getattr(tab, "azm_value", tab.azm))
            if tab.amp_mode == 2 and not hasattr(tab, "azm_value"):
                raise ValueError("vjp needs the ksq tables or streaming tables made from them")
            ampbar = per_bin * (true_azm / float(tab.azm))
            azmbar = float((amps[0] * per_bin).sum() / float(tab.azm))
            return xibar, [ampbar], azmbar, float(ybar.sum(dtype=torch.float64))
        # general grids / product spaces
        h = ybar
        for pl, g in zip(reversed(self.plans), reversed(self.grids)):
            h = pl.hartley(h, convention=convention, epi=PointOp(scale=1.0 / g.total_volume))
        full = tab.amp if tab.amp_mode == 2 else None
        if full is None:
            raise ValueError("vjp: general grids use amp_mode-2 tables")
        true_azm = float(getattr(tab, "azm_value", tab.azm))
        xibar = h * full * float(tab.azm)
        u = (xi.to(torch.float64) * h.to(torch.float64))
        ea = None
        exps = [g.expand(a, torch.float64) for a, g in zip(amps, self.grids)]
        azmbar = None
        ampbars = []
        for i, (g, ax) in enumerate(zip(self.grids, self.axes)):
            w = true_azm * u
            for j, (e, axj) in enumerate(zip(exps, self.axes)):
                if j == i:
                    continue
                shp = [1] * len(self.shape)
                for k, axk in enumerate(axj):
                    shp[axk] = self.shape[axk]
                w = w * e.reshape(shp)
            red = tuple(k for k in range(len(self.shape)) if k not in ax)
            w = w.sum(dim=red) if red else w
            ampbars.append(torch.zeros(amps[i].numel(), dtype=torch.float64, device=self.device)
                           .index_add_(0, g.pindex.to(self.device).reshape(-1), w.reshape(-1)))
            if azmbar is None:
                azmbar = float((ampbars[0] * amps[0]).sum() / true_azm)
        return xibar, ampbars, azmbar, float(ybar.sum(dtype=torch.float64))

    def __call__(self, p, convention=None):
        """Full model evaluation.  numpy in -> numpy out (host buffers, copies included);
        torch device tensors in -> device tensor out."""
        xi = p[self.prefix + "xi"]
        host = not isinstance(xi, torch.Tensor)
        tab = self.tables(p)
        y = self.apply_xi(self._as_dev(xi, self.dtype).contiguous(), tab, convention=convention)
        return y.cpu().numpy() if host else y


def SimpleCorrelatedField(shape, distances, offset_mean, offset_std, fluctuations, flexibility, asperity,
                          loglogavgslope, prefix="", dtype=torch.float64, non_parametric_kind="power"):
    """Single-space correlated field (``src/library/correlated_fields_simple.py:38-170``;
    same argument order).  The numpy stack parametrises the *power* spectrum
    (``test/test_re/test_correlated_field.py:150``)."""
    mk = CorrelatedFieldMaker(prefix, dtype=dtype)
    mk.set_amplitude_total_offset(offset_mean, offset_std)
    mk.add_fluctuations(shape, distances, fluctuations, loglogavgslope, flexibility, asperity,
                        non_parametric_kind=non_parametric_kind)
    return mk.finalize()
