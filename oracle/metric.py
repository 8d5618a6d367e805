"""Oracle: the MGVI/geoVI metric that the CG inverts, and the linear sample draw.

Test infrastructure.  Restates, for a correlated field with a pointwise link and
a diagonal likelihood,
* ``ham_metric(p, t) = lh.metric(p, t) + t`` (``/root/reference/src/re/evi.py:108-109``),
* ``LikelihoodWithModel.metric`` = ``J^T . lh.metric(f(p), J t)``
  (``src/re/likelihood.py:568-576``),
* ``Gaussian.metric`` = ``noise_cov_inv(t)`` (``src/re/likelihood_impl.py:136-137``),
  ``Poissonian.metric`` = ``t / primals`` (``:248-249``),
* ``draw_linear_residual`` (``src/re/evi.py:83-144``): ``j = J^T lsm(white) + white_prior``,
  ``x0 = white_prior``, solved with CG on ``ham_metric``,
* numpy-stack twin: ``SamplingEnabler`` (``src/operators/sampling_enabler.py:59-86``) with
  ``SandwichOperator.make(J)`` (``src/operators/sandwich_operator.py:42-76``).

For the xi-block (the N-sized hot path) with fixed amplitude tables:
``M t = A H diag(w) H A t + t`` with ``A = diag(azm * outer(amps)[pindex])``,
``H`` = chained ``(1/V) Hartley`` and ``w = f'(s)^2 / sigma^2`` (Gaussian) or
``w = lambda = exp(s)`` (Poisson with exp link).
"""

import numpy as np

from . import transforms as _tr


def gaussian_weight(noise_cov_inv, link_derivative=None):
    """``w`` for a Gaussian likelihood: ``f'(s)^2 * N^-1`` (``likelihood_impl.py:136-137``)."""
    w = np.asarray(noise_cov_inv, dtype=np.float64)
    return w if link_derivative is None else w * np.asarray(link_derivative) ** 2


def poisson_exp_weight(signal):
    """``w`` for ``Poissonian`` with ``lambda = exp(s)``: ``J_f^T (1/lambda) J_f`` with
    ``J_f = diag(lambda) J_s`` gives ``diag(lambda)`` (``likelihood_impl.py:248-249``)."""
    return np.exp(np.asarray(signal, dtype=np.float64))


def cf_metric_xi(cf, amps, azm, w, convention=_tr.NON_CANONICAL):
    """Return ``mat(t) = A H w H A t + t`` (``evi.py:108-109`` on the xi block)."""
    A = azm * cf.outer_amplitude(amps)

    def H(x):
        return cf.harmonic_transform(x, convention)

    def Ht(x):
        out = x
        for g, ax in reversed(list(zip(cf.grids, cf.axes))):
            out = (1.0 / g.total_volume) * _tr.hartley(out, axes=ax, convention=convention)
        return out

    def mat(t):
        return A * Ht(w * H(A * t)) + t

    return mat


def linear_sample_rhs(cf, amps, azm, w, white_data, white_prior, convention=_tr.NON_CANONICAL):
    """``nll_smpl + prr_smpl`` of ``draw_linear_residual`` (``evi.py:114-125``):
    ``j = J^T sqrt(w) eta + x0`` with ``x0 = white_prior``."""
    A = azm * cf.outer_amplitude(amps)
    h = np.sqrt(w) * white_data
    for g, ax in reversed(list(zip(cf.grids, cf.axes))):
        h = (1.0 / g.total_volume) * _tr.hartley(h, axes=ax, convention=convention)
    return A * h + white_prior
