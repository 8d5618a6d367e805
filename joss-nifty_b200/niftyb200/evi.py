"""MGVI / geoVI building blocks on a correlated-field model with a Gaussian likelihood.

Mirrors, on flat device latent vectors (:class:`niftyb200.latent.LatentLayout`):
* ``Gaussian`` + ``LikelihoodWithModel`` (``/root/reference/src/re/likelihood_impl.py:96-143``,
  ``src/re/likelihood.py:520-640``): energy, ``transformation``, ``left_sqrt_metric`` = ``J^T sqrt(N^-1) .``,
  ``right_sqrt_metric`` = ``sqrt(N^-1) J .``, ``metric``;
* ``_StandardHamiltonian`` and the sample-averaged ``_kl_vg`` / ``_kl_met`` (``src/re/optimize_kl.py:60-120``);
* the geoVI update ``nonlinearly_update_residual`` with ``_residual_vg`` / ``_metric`` / ``_sampnorm``
  (``src/re/evi.py:147-266``), minimised with :func:`niftyb200.optimize.newton_cg`.
Every Jacobian / adjoint apply is one fused transform plus the O(n_bins) chain kernels (``jvp_latent`` /
``vjp_latent``); the host only orchestrates.
"""

import numpy as np
import torch

from .latent import LatentLayout, LatentMetric, draw_linear_residual
from .optimize import OptimizeResults, newton_cg
from .transforms import vdot_device


class GaussianCFModel:
    """``data = cf(latent) + noise``, ``noise ~ N(0, 1/noise_cov_inv)`` (diagonal)."""

    def __init__(self, cf, data, noise_cov_inv, convention=None):
        self.cf, self.convention = cf, convention
        self.layout = LatentLayout(cf)
        self.data = data.to(device=cf.device, dtype=cf.dtype).contiguous()
        self.w = noise_cov_inv.to(device=cf.device, dtype=cf.dtype).contiguous()
        self.sqrt_w = torch.sqrt(self.w)
        self._cache = []          # (key, p, tab) of the last expansion points

    # ---- expansion points -----------------------------------------------------------------
    def at(self, x):
        """(parameter dict of views, amplitude tables + chain context) at the flat latent ``x``; cached."""
        key = (x.data_ptr(), x._version, x.numel())
        for k, p, tab in self._cache:
            if k == key:
                return p, tab
        p = self.layout.unflatten(x)
        tab = self.cf.tables(p)
        self._cache = [(key, p, tab)] + self._cache[:3]
        return p, tab

    def signal(self, x, out=None):
        p, tab = self.at(x)
        return self.cf.apply_xi(p[self.cf.prefix + "xi"].contiguous(), tab, out=out, convention=self.convention)

    def jac(self, x, t, mul=None):
        p, tab = self.at(x)
        return self.cf.jvp_latent(self.layout.unflatten(t), p, convention=self.convention, tab=tab, mul=mul)

    def jac_adjoint(self, x, ybar, add=None):
        p, tab = self.at(x)
        adj = self.cf.vjp_latent(ybar, p, convention=self.convention, tab=tab)
        adj.pop("offset_mean", None)
        out = self.layout.flatten(adj)
        return out if add is None else out.add_(add)

    # ---- Likelihood interface (likelihood.py:520-640) -----------------------------------------
    def transformation(self, x):
        """``sqrt(N^-1) f(x)``: the coordinates in which the likelihood metric is the identity (likelihood_impl.py:139-143)."""
        return self.sqrt_w * self.signal(x)

    def left_sqrt_metric(self, x, y):
        return self.jac_adjoint(x, (self.sqrt_w * y).contiguous())

    def right_sqrt_metric(self, x, t):
        return self.jac(x, t, mul=self.sqrt_w)

    def metric_at(self, x):
        """``ham_metric`` at ``x`` as an operator for the CG."""
        p, _ = self.at(x)
        return LatentMetric(self.cf, p, self.w, convention=self.convention)

    # ---- Hamiltonian (optimize_kl.py:60-81) --------------------------------------------------------
    def energy_and_grad(self, x):
        """``0.5 (d - s)^T N^-1 (d - s) + 0.5 x.x`` and its gradient ``-J^T N^-1 (d - s) + x``."""
        res = self.data - self.signal(x)
        wres = self.w * res
        e = 0.5 * float(vdot_device(res, wres).cpu()) + 0.5 * float(vdot_device(x, x).cpu())
        g = self.jac_adjoint(x, wres.neg_().contiguous(), add=x)
        return e, g

    # ---- sample-averaged KL (optimize_kl.py:84-120) ------------------------------------------------
    def kl_energy_and_grad(self, samples):
        def fun(pos):
            if len(samples) == 0:
                return self.energy_and_grad(pos)
            e, g = 0.0, None
            for s in samples:
                es, gs = self.energy_and_grad(pos + s)
                e += es
                g = gs if g is None else g.add_(gs)
            return e / len(samples), g.div_(len(samples))
        return fun

    def kl_metric(self, samples):
        model = self

        class _At:
            def at(self, pos):
                ms = [model.metric_at(pos + s) for s in samples] if len(samples) else [model.metric_at(pos)]

                def apply(t, out=None):
                    acc = None
                    for m in ms:
                        q = m(t)
                        acc = q if acc is None else acc.add_(q)
                    return acc.div_(len(ms))
                return apply
        return _At()


# ---------------------------------------------------------------- geoVI (evi.py:147-266)
def _residual_vg(model, e, trafo_at_p, ms_at_p, x):
    """``_residual_vg`` (evi.py:160-174)."""
    t = model.transformation(x) - trafo_at_p
    g = x - e + model.left_sqrt_metric(e, t)
    r = ms_at_p - g
    res = 0.5 * float(vdot_device(r, r).cpu())
    ngrad = r + model.left_sqrt_metric(x, model.right_sqrt_metric(e, r))
    return res, -ngrad


def _geovi_metric(model, e, primals, tangents):
    """``_metric`` (evi.py:176-183)."""
    tm = model.left_sqrt_metric(e, model.right_sqrt_metric(primals, tangents)) + tangents
    return model.left_sqrt_metric(primals, model.right_sqrt_metric(e, tm)) + tm


def _sampnorm(model, e, natgrad):
    """``_sampnorm`` (evi.py:185-190)."""
    fpp = model.right_sqrt_metric(e, natgrad)
    return float(np.sqrt(float(vdot_device(natgrad, natgrad).cpu()) + float(vdot_device(fpp, fpp).cpu())))


def nonlinearly_update_residual(model, pos, residual_sample, white_data, white_prior, metric_sample_sign=1.0, *,
                                minimize=newton_cg, minimize_kwargs=None):
    """``nonlinearly_update_residual`` (evi.py:201-266): the geoVI correction of an MGVI residual sample.
    ``white_data`` / ``white_prior`` replace the PRNG key of the metric sample (``draw_linear_residual`` with
    ``from_inverse=False``, :155-158).  Returns ``(updated residual sample, OptimizeResults)``."""
    minimize_kwargs = dict(minimize_kwargs or {})
    sample = pos + residual_sample                                               # :222
    metric_sample, _ = draw_linear_residual(model.metric_at(pos), white_data, white_prior, from_inverse=False)   # :225-227
    metric_sample = metric_sample * metric_sample_sign
    skip = isinstance(minimize_kwargs.get("maxiter", None), int) and minimize_kwargs["maxiter"] == 0
    if not skip:
        trafo_at_p = model.transformation(pos)                                   # :236
        opt_state = minimize(lambda x: _residual_vg(model, pos, trafo_at_p, metric_sample, x), sample,
                             lambda x, t: _geovi_metric(model, pos, x, t),
                             custom_gradnorm=lambda v: _sampnorm(model, pos, v), **minimize_kwargs)
    else:
        opt_state = OptimizeResults(sample, True, 0, None, None)
    sample = opt_state.x - pos                                                   # :259-261
    return sample, opt_state._replace(x=None, jac=None)
