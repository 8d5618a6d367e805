"""Oracle: the O(n_bins) amplitude (power-spectrum) chain of the correlated field.

Test infrastructure.  NumPy restatement of the JAX stack, which cannot be
imported here (jax absent):
* ``normal_prior`` / ``lognormal_prior`` / ``lognormal_moments``
  (``/root/reference/src/re/num/stats_distributions.py:41-88``),
* ``integrated_wiener_process`` (``src/re/gauss_markov.py:105-121``),
* ``_remove_slope`` (``src/re/correlated_field.py:145-147``),
* ``non_parametric_amplitude.correlate`` (``:277-304``),
* ``matern_amplitude.correlate`` (``:186-213``).
Pinned against the importable numpy stack (``src/library/correlated_fields.py``
``_Amplitude`` ``:276-385``) through ``tests/golden`` exactly like
``test/test_re/test_correlated_field.py:120-167``.
"""

import numpy as np


def lognormal_moments(mean, std):
    """``stats_distributions.py:57-68``."""
    if mean <= 0.0:
        raise ValueError(f"`mean` must be greater zero; got {mean!r}")
    if std <= 0.0:
        raise ValueError(f"`std` must be greater zero; got {std!r}")
    logstd = np.sqrt(np.log1p((std / mean) ** 2))
    logmean = np.log(mean) - 0.5 * logstd ** 2
    return logmean, logstd


def normal_prior(mean, std):
    """``stats_distributions.py:41-45``: xi -> mean + std*xi."""
    return lambda xi: mean + std * xi


def lognormal_prior(mean, std):
    """``stats_distributions.py:75-88``: xi -> exp(logmean + logstd*xi)."""
    lm, ls = lognormal_moments(mean, std)
    return lambda xi: np.exp(lm + ls * xi)


def integrated_wiener_process(xi, x0, sigma, dt, asperity=None):
    """``gauss_markov.py:105-121``; ``xi`` has shape (len(dt), 2)."""
    asperity = 0.0 if asperity is None else asperity
    xi = np.asarray(xi, dtype=np.float64)
    dt = np.ones(xi.shape[0]) * dt if np.isscalar(dt) else np.asarray(dt, dtype=np.float64)
    res = (sigma * np.sqrt(dt))[:, np.newaxis] * xi
    res[:, 0] *= np.sqrt(dt ** 2 / 12.0 + asperity)
    res[:, 0] += 0.5 * dt * res[:, 1]
    res = np.concatenate((np.asarray(x0, dtype=np.float64)[np.newaxis, ...], res), axis=0)
    res[:, 1] = np.cumsum(res[:, 1])
    res[1:, 0] += dt * res[:-1, 1]
    res[:, 0] = np.cumsum(res[:, 0])
    return res


def remove_slope(rel_log_mode_dist, x):
    """``re/correlated_field.py:145-147``."""
    sc = rel_log_mode_dist / rel_log_mode_dist[-1]
    return x - x[-1] * sc


def non_parametric_amplitude(grid, fluctuations, loglogavgslope, flexibility=None,
                             asperity=None, spectrum_xi=None, kind="amplitude"):
    """``non_parametric_amplitude.correlate`` (``re/correlated_field.py:277-304``).

    ``fluctuations`` ... ``asperity`` are the already prior-transformed scalar
    values (or None); ``spectrum_xi`` the white (nb-2, 2) excitations of the IWP.
    """
    totvol = grid.total_volume
    rel_log_mode_len = grid.relative_log_mode_lengths
    mode_multiplicity = grid.mode_multiplicity
    log_vol = grid.log_volume

    ln_spectrum = loglogavgslope * rel_log_mode_len
    if flexibility is not None and log_vol.size > 0:
        twolog = integrated_wiener_process(spectrum_xi, np.zeros((2,)), flexibility, log_vol,
                                           asperity=asperity)
        twolog = np.concatenate((np.zeros((1,)), twolog[:, 0]))
        ln_spectrum = ln_spectrum + remove_slope(rel_log_mode_len, twolog)

    spectrum = np.exp(ln_spectrum)
    if kind.lower() == "amplitude":
        norm = np.sqrt(np.sum(mode_multiplicity[1:] * spectrum[1:] ** 2))
        norm /= np.sqrt(totvol)
        amplitude = fluctuations * (np.sqrt(totvol) / norm) * spectrum
    elif kind.lower() == "power":
        norm = np.sqrt(np.sum(mode_multiplicity[1:] * spectrum[1:]))
        norm /= np.sqrt(totvol)
        amplitude = fluctuations * (np.sqrt(totvol) / norm) * np.sqrt(spectrum)
    else:
        raise ValueError(f"invalid kind specified {kind!r}")
    amplitude = amplitude.copy()
    amplitude[0] = totvol
    return amplitude


def matern_amplitude(grid, scale, cutoff, loglogslope, renormalize_amplitude=False,
                     kind="amplitude"):
    """``matern_amplitude.correlate`` (``re/correlated_field.py:186-213``)."""
    totvol = grid.total_volume
    mode_lengths = grid.mode_lengths
    mode_multiplicity = grid.mode_multiplicity
    ln_spectrum = 0.25 * loglogslope * np.log1p((mode_lengths / cutoff) ** 2)
    spectrum = np.exp(ln_spectrum)
    norm = 1.0
    if renormalize_amplitude:
        if kind.lower() == "amplitude":
            norm = np.sqrt(np.sum(mode_multiplicity[1:] * spectrum[1:] ** 4))
        elif kind.lower() == "power":
            norm = np.sqrt(np.sum(mode_multiplicity[1:] * spectrum[1:] ** 2))
        norm /= np.sqrt(totvol)
    spectrum = scale * (np.sqrt(totvol) / norm) * spectrum
    spectrum[0] = totvol
    if kind.lower() == "power":
        spectrum = np.sqrt(spectrum)
    elif kind.lower() != "amplitude":
        raise ValueError(f"invalid kind specified {kind!r}")
    return spectrum
