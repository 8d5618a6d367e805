"""The metric over the FULL latent vector of a correlated-field model and the MGVI sample draw.

Reference: ``LikelihoodWithModel.metric`` = ``J^T lh.metric(f(p), J t)`` (``/root/reference/src/re/likelihood.py:568-576``),
``ham_metric = lh.metric + identity`` and ``draw_linear_residual`` (``src/re/evi.py:83-144``); numpy stack
``SamplingEnabler`` on ``SandwichOperator.make(J)`` (``src/operators/sampling_enabler.py:59-86``).  The MGVI sample
lives on ``xi`` AND on the spectrum / zero-mode leaves: ``J`` is the Jacobian of the whole model with respect to
every latent leaf, not only its ``xi`` block (:class:`niftyb200.metric.CFMetric`).

One apply = ``jvp_latent`` (amplitude-chain JVP kernel + one fused transform, the likelihood weight on its last
store) followed by ``vjp_latent`` (one fused transform with the binned reduction + the chain's VJP kernel).
The conjugate-gradient vectors are single flat device tensors (``xi`` first, then the small leaves), so the
fused CG vector kernels run on them unchanged.
"""

import numpy as np
import torch

from .cg import CGResults, static_cg, _static_cg, _cg  # noqa: F401


class LatentLayout:
    """Flat layout of the latent leaves of a :class:`CorrelatedField`: ``xi`` first (so that it stays 16-byte
    aligned and is a plain view), then the other leaves in sorted order."""

    def __init__(self, cf):
        self.cf = cf
        self.keys = [cf.prefix + "xi"] + sorted(k for k in cf.domain if k != cf.prefix + "xi")
        self.shapes = {k: tuple(cf.domain[k]) for k in self.keys}
        self.offsets, n = {}, 0
        for k in self.keys:
            self.offsets[k] = n
            n += int(np.prod(self.shapes[k], dtype=np.int64))
        self.size = n

    def flatten(self, d, out=None):
        cf = self.cf
        out = torch.zeros(self.size, dtype=cf.dtype, device=cf.device) if out is None else out
        for k in self.keys:
            v = d.get(k)
            if v is None:
                continue
            v = v if isinstance(v, torch.Tensor) else torch.as_tensor(np.asarray(v))
            o = self.offsets[k]
            out[o:o + v.numel()] = v.to(device=cf.device, dtype=cf.dtype).reshape(-1)
        return out

    def unflatten(self, t):
        """dict of VIEWS into the flat tensor ``t``."""
        out = {}
        for k in self.keys:
            o, shp = self.offsets[k], self.shapes[k]
            n = int(np.prod(shp, dtype=np.int64))
            out[k] = t[o:o + n].reshape(shp)
        return out


class LatentMetric:
    """``M t = J^T diag(w) J t + t`` over the full latent vector at the expansion point ``p``
    (``ham_metric``, evi.py:108-109).  ``weight``: the diagonal likelihood metric in position space
    (Gaussian ``noise_cov_inv``, Poissonian ``1 / lambda`` composed with the link, see :mod:`niftyb200.metric`)."""

    def __init__(self, cf, p, weight, convention=None):
        self.cf, self.p, self.convention = cf, p, convention
        self.layout = LatentLayout(cf)
        self.tab = cf.tables(p)
        self.w = weight.to(device=cf.device, dtype=cf.dtype).contiguous()
        self._u = torch.empty(cf.shape, dtype=cf.dtype, device=cf.device)
        self.shape, self.dtype, self.device = (self.layout.size,), cf.dtype, cf.device
        self.napply = 0

    def jacobian(self, t, mul=None, out=None):
        """``J t`` (position space) for a flat latent tangent."""
        return self.cf.jvp_latent(self.layout.unflatten(t), self.p, convention=self.convention, tab=self.tab, mul=mul, out=out)

    def jacobian_adjoint(self, ybar, out=None, add=None):
        """``J^T ybar`` as a flat latent vector (``+ add``)."""
        adj = self.cf.vjp_latent(ybar, self.p, convention=self.convention, tab=self.tab)
        adj.pop("offset_mean", None)
        out = self.layout.flatten(adj, out=out)
        if add is not None:
            out.add_(add)
        return out

    def __call__(self, t, out=None):
        self.napply += 1
        u = self.jacobian(t, mul=self.w, out=self._u)
        if out is not None and out.data_ptr() == t.data_ptr():
            raise ValueError("LatentMetric: out must not alias the input")
        return self.jacobian_adjoint(u, out=out, add=t)

    def sample_rhs(self, white_data, white_prior):
        """``nll_smpl + prr_smpl`` (evi.py:114-125): ``J^T sqrt(w) eta + xi_prior`` (flat)."""
        h = torch.sqrt(self.w) * white_data.to(device=self.device, dtype=self.dtype)
        return self.jacobian_adjoint(h.contiguous(), add=white_prior)


def draw_linear_residual(metric, white_data, white_prior, *, from_inverse=True, cg=static_cg, cg_name=None,
                         cg_kwargs=None, _raise_nonposdef=False):
    """``draw_linear_residual`` (evi.py:83-144) on a :class:`LatentMetric` (or :class:`CFMetric`): the white
    excitations ``white_data`` (position space) and ``white_prior`` (latent space) take the place of the two
    PRNG sub-keys (parity is on the same inputs: JAX's generator cannot be matched).  Returns ``(sample, info)``;
    raises ``ValueError("conjugate gradient failed")`` when ``info < 0`` (:139-142)."""
    smpl = metric.sample_rhs(white_data, white_prior)                      # :114-125
    info = 0
    if from_inverse:
        kw = dict(cg_kwargs or {})
        if cg in (static_cg,):
            smpl, info = cg(metric, smpl, x0=white_prior, name=cg_name, **kw)           # :131-138
        else:
            smpl, info = cg(metric, smpl, x0=white_prior, name=cg_name, _raise_nonposdef=_raise_nonposdef, **kw)
        if info is not None and info < 0:
            raise ValueError("conjugate gradient failed")
    return smpl, info
