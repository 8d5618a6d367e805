"""The operator the conjugate gradient inverts: the MGVI / geoVI metric of a correlated field.

``M t = J^T N^-1 J t + t`` (reference: ``ham_metric`` ``/root/reference/src/re/evi.py:108-109``;
``LikelihoodWithModel.metric`` ``src/re/likelihood.py:568-576``; numpy stack
``SamplingEnabler`` ``src/operators/sampling_enabler.py:59`` with ``SandwichOperator.make``
``src/operators/sandwich_operator.py:42-76``).  For the xi block of a correlated field with a
point-wise link and a diagonal likelihood this is

    M t = A H diag(w) H A t + t,   A = diag(azm * amp[pindex]),  H = prod_g (1/V_g) Hartley_g,

with ``w = f'(s)^2 / sigma^2`` (Gaussian, ``src/re/likelihood_impl.py:136-137``) or
``w = lambda = exp(s)`` (Poissonian with exp link, ``:248-249``).

One metric apply = two fused 3-pass transforms: the weight multiply rides on the last store
of the forward transform, the amplitude multiply, the ``+ t`` and the CG curvature ``t.(M t)``
on the last store of the adjoint transform.
"""

import os

import torch

from .transforms import PointOp


def gaussian_weight(noise_cov_inv, link_derivative=None):
    """``w = f'(s)^2 * N^-1`` (likelihood_impl.py:136-137)."""
    return noise_cov_inv if link_derivative is None else noise_cov_inv * link_derivative ** 2


def poisson_exp_weight(signal):
    """``w = exp(s)`` for ``Poissonian`` with ``lambda = exp(s)`` (likelihood_impl.py:248-249)."""
    return torch.exp(signal)


class CFMetric:
    """``mat(t) = A H w H A t + t`` on device tensors."""

    def __init__(self, cf, tables, weight, convention=None, stream_amplitude=True):
        self.cf, self.convention = cf, convention
        # the expansion point is fixed for the whole solve: resolve the amplitude gather once
        self.tab = cf.streaming_tables(tables) if stream_amplitude else tables
        # slab-decomposed fields: latent vectors and position-space fields have different local shapes
        self.latent_shape = tuple(getattr(cf, "latent_shape", cf.shape))
        self.position_shape = tuple(getattr(cf, "position_shape", cf.shape))
        self.comm = getattr(cf, "comm", None)
        if tuple(weight.shape) != self.position_shape:
            raise ValueError("weight must have the (local) position-space shape")
        self.w = weight.to(device=cf.device, dtype=cf.dtype).contiguous()
        self._u = torch.empty(self.position_shape, dtype=cf.dtype, device=cf.device)
        self._curv = torch.zeros((), dtype=torch.float64, device=cf.device)
        self.shape, self.dtype, self.device = self.latent_shape, cf.dtype, cf.device
        self.napply = 0
        # apply_with_curvature(curv_out=..., pre=(r, device beta)) makes no host read-back: static_cg may keep
        # its whole loop on the device
        self.device_resident_cg = True
        # the curvature is reduced exactly once per apply, by this library, into `curv_out`: the scalar step of the
        # device-resident CG may ride on that launch (nb2_set_cg_fusion)
        self.fuses_cg_decide = True
        # the CG loops hand their pending direction update d <- r + beta d to apply_with_curvature(pre=...)
        # (NIFTYB200_FUSED_AXPBY=0: keep the separate axpby kernel -- A/B measurements)
        self.fused_direction_update = (bool(getattr(cf, "fuses_direction_update", False))
                                       and os.environ.get("NIFTYB200_FUSED_AXPBY", "1") != "0")

    def forward_weighted(self, t, out, pre=None):
        """``w * H A t`` (offset excluded: this is the Jacobian); ``pre``: see ``CorrelatedField.apply_xi``."""
        if pre is not None:
            return self.cf.apply_xi(t, self.tab, out=out, convention=self.convention, offset=0.0, mul=self.w, pre=pre)
        return self.cf.apply_xi(t, self.tab, out=out, convention=self.convention, offset=0.0, mul=self.w)

    def __call__(self, t, out=None):
        self.napply += 1
        u = self.forward_weighted(t, self._u)
        return self.cf.adjoint_xi(u, self.tab, out=out, convention=self.convention, add=t)

    def apply_with_curvature(self, t, out=None, pre=None, curv_out=None):
        """Returns ``(M t, t.(M t))``; the dot product is fused into the last store.  The
        curvature is a 0-d fp64 device tensor (``curv_out``: where to put it, e.g. a slot of the
        device-resident CG state).  ``pre = (r, beta)``: ``t <- r + beta t`` in place first (fused into the
        first pass; only when :attr:`fused_direction_update`); ``beta`` may be a 1-element fp64 device tensor."""
        self.napply += 1
        curv = self._curv if curv_out is None else curv_out
        u = self.forward_weighted(t, self._u, pre)
        q = self.cf.adjoint_xi(u, self.tab, out=out, convention=self.convention, add=t, dot=t, dot_out=curv)
        return q, curv

    def sample_rhs(self, white_data, white_prior, out=None):
        """``j = J^T sqrt(w) eta + x0`` of ``draw_linear_residual`` (src/re/evi.py:114-125)."""
        h = torch.sqrt(self.w) * white_data
        return self.cf.adjoint_xi(h, self.tab, out=out, convention=self.convention, add=white_prior)


class SampledCFMetric:
    """The sample-averaged metric the KL minimisation inverts:

        M t = (1/S) sum_s (A_s H w_s H A_s t) + t

    i.e. ``_kl_met`` (``src/re/optimize_kl.py:103-120``: ``ham.metric`` mapped over the samples and averaged with
    ``reduce``) / ``SampledKLEnergyClass.apply_metric`` (``src/minimization/kl_energies.py:340-345``:
    ``sample_list.average``) on the xi block.  Every sample is an expansion point of its own: amplitude tables
    ``A_s`` and likelihood weight ``w_s``.

    Per sample: one forward and one adjoint fused transform; the adjoint's epilogue scales by ``1/S`` and adds
    the running sum in place, so the average costs no sweep of its own, and the curvature ``t.(M t)`` rides on
    the last sample's last store.  ``comm`` (a :class:`niftyb200.distributed.Comm`-like object with
    ``allreduce_``, ``size``): the samples are sharded over the ranks -- the reference's MPI sample parallelism
    (``kl_energies.py:136-141``, ``utilities.py:331-390``) -- and the partial sums are all-reduced (one
    N-vector per apply); ``n_total`` is then the global number of samples."""

    def __init__(self, cf, tables_list, weights, convention=None, comm=None, n_total=None):
        if len(tables_list) != len(weights):
            raise ValueError("one weight per sample")
        self.cf, self.convention, self.sample_comm = cf, convention, comm
        self.tabs = [cf.streaming_tables(t) for t in tables_list]
        self.ws = [w.to(device=cf.device, dtype=cf.dtype).contiguous() for w in weights]
        self.n_local = len(self.ws)
        self.n_total = int(n_total) if n_total is not None else self.n_local * (comm.size if comm is not None else 1)
        if self.n_total < 1:
            raise ValueError("at least one sample")
        self.shape, self.dtype, self.device = tuple(cf.shape), cf.dtype, cf.device
        self._u = torch.empty(self.shape, dtype=cf.dtype, device=cf.device)
        # the running sum alternates between `out` and this buffer: a transform uses its output buffer for the
        # intermediate passes, so the addend must never alias the output
        self._acc = torch.empty(self.shape, dtype=cf.dtype, device=cf.device) if self.n_local > 1 else None
        self._curv = torch.zeros((), dtype=torch.float64, device=cf.device)
        self.napply = 0
        # with samples on other ranks the curvature needs the all-reduced result: computed afterwards, on the host side
        self.device_resident_cg = comm is None and self.n_local > 0
        self.fuses_cg_decide = self.device_resident_cg

    def __call__(self, t, out=None):
        return self._apply(t, out, None, None)[0]

    def apply_with_curvature(self, t, out=None, pre=None, curv_out=None):
        """``(M t, t.(M t))``; see :meth:`CFMetric.apply_with_curvature`."""
        if pre is not None:
            raise ValueError("SampledCFMetric does not carry the fused direction update")
        return self._apply(t, out, t, self._curv if curv_out is None else curv_out)

    def _apply(self, t, out, dot, curv):
        self.napply += 1
        cf, S = self.cf, self.n_total
        out = torch.empty_like(t) if out is None else out
        local = self.sample_comm is None
        prev = None
        for s, (tab, w) in enumerate(zip(self.tabs, self.ws)):
            u = cf.apply_xi(t, tab, out=self._u, convention=self.convention, offset=0.0, mul=w)
            last = s == self.n_local - 1
            dst = out if (self.n_local - 1 - s) % 2 == 0 else self._acc       # the last sample writes `out`
            # first sample: + t (the prior's identity) when no other rank contributes; later: + running sum
            add = (t if local else None) if s == 0 else prev
            prev = dst
            epi = PointOp(tab.amp_mode, tab.amp, tab.azm * cf.dvol / S, add=add, amp_sym=getattr(tab, "sym", False),
                          dot=dot if (last and local) else None, dot_out=curv if (last and local) else None)
            if cf.single:
                cf.plans[0].hartley(u, out=dst, convention=self.convention, epi=epi)
            else:
                cur = cf.plans[-1].hartley(u, out=u, convention=self.convention)
                for pl in reversed(cf.plans[1:-1]):
                    cur = pl.hartley(cur, out=cur, convention=self.convention)
                cf.plans[0].hartley(cur, out=dst, convention=self.convention, epi=epi)
        if not local:
            if self.n_local == 0:
                out.zero_()
            self.sample_comm.allreduce_(out)
            out.add_(t)
            if dot is not None:
                from .transforms import vdot_device
                curv.copy_(vdot_device(dot, out))
        return out, curv
