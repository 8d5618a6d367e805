"""Oracle: conjugate-gradient solvers of both reference stacks (test infrastructure).

NumPy restatements, fp64, of
* ``ConjugateGradient.__call__`` (``/root/reference/src/minimization/conjugate_gradient.py:48-126``)
  with ``QuadraticEnergy`` (``src/minimization/quadratic_energy.py:25-78``) and the
  controllers ``GradientNormController`` / ``AbsDeltaEnergyController``
  (``src/minimization/iteration_controllers.py:148-221,356-423``);
* ``_cg`` (``src/re/conjugate_gradient.py:77-216``) and ``_static_cg`` (``:219-389``).

Vectors are plain ndarrays; ``vdot`` is ``np.vdot`` on the flattened arrays
(the reference's tree ``vdot`` sums per-leaf ``jnp.vdot``: ``re/tree_math/vector_math.py:218-220``).
"""

from collections import namedtuple

import numpy as np

CONVERGED, CONTINUE, ERROR = 0, 1, 2
N_RESET = 20


def _vdot(a, b):
    """``ducc_dispatch.vdot`` (``src/ducc_dispatch.py:81-86``): ``ducc0.misc.vdot`` accumulates in extended
    precision whatever the input precision (the SciPy fallback warns about exactly that for fp32,
    ``:53-58``), so fp32 operands are accumulated in fp64 here."""
    a, b = np.ravel(a), np.ravel(b)
    if a.dtype == np.float32 or b.dtype == np.float32:
        a, b = a.astype(np.float64), b.astype(np.float64)
    return float(np.vdot(a, b).real)


# ------------------------------------------------------------------ numpy stack
class QuadraticEnergy:
    """``quadratic_energy.py:25-78``."""

    def __init__(self, position, A, b, _grad=None):
        self.position, self._A, self._b = position, A, b
        if _grad is not None:
            self.gradient = _grad
            Ax = _grad if b is None else _grad + b
        else:
            Ax = A(position)
            self.gradient = Ax if b is None else Ax - b
        self.value = 0.5 * _vdot(position, Ax)
        if b is not None:
            self.value -= _vdot(b, position)
        self._gradnorm = None

    def at(self, position):
        return QuadraticEnergy(position, self._A, self._b)

    def at_with_grad(self, position, grad):
        return QuadraticEnergy(position, self._A, self._b, grad)

    @property
    def gradient_norm(self):
        if self._gradnorm is None:
            self._gradnorm = float(np.sqrt(_vdot(self.gradient, self.gradient)))
        return self._gradnorm

    def apply_metric(self, x):
        return self._A(x)


class GradientNormController:
    """``iteration_controllers.py:148-221``."""

    def __init__(self, tol_abs_gradnorm=None, tol_rel_gradnorm=None, convergence_level=1,
                 iteration_limit=None):
        self._tol_abs_gradnorm, self._tol_rel_gradnorm = tol_abs_gradnorm, tol_rel_gradnorm
        self._convergence_level, self._iteration_limit = convergence_level, iteration_limit

    def start(self, energy):
        self._itcount, self._ccount = -1, 0
        if self._tol_rel_gradnorm is not None:
            self._tol_rel_gradnorm_now = self._tol_rel_gradnorm * energy.gradient_norm
        return self.check(energy)

    def check(self, energy):
        self._itcount += 1
        inclvl = False
        if self._tol_abs_gradnorm is not None and energy.gradient_norm <= self._tol_abs_gradnorm:
            inclvl = True
        if self._tol_rel_gradnorm is not None and energy.gradient_norm <= self._tol_rel_gradnorm_now:
            inclvl = True
        self._ccount = self._ccount + 1 if inclvl else max(0, self._ccount - 1)
        if self._iteration_limit is not None and self._itcount >= self._iteration_limit:
            return CONVERGED
        if self._ccount >= self._convergence_level:
            return CONVERGED
        return CONTINUE


class AbsDeltaEnergyController:
    """``iteration_controllers.py:356-423``."""

    def __init__(self, deltaE, convergence_level=1, iteration_limit=None):
        self._deltaE, self._convergence_level, self._iteration_limit = deltaE, convergence_level, iteration_limit

    def start(self, energy):
        self._itcount, self._ccount, self._Eold = -1, 0, 0.0
        return self.check(energy)

    def check(self, energy):
        self._itcount += 1
        Eval = energy.value
        diff = abs(self._Eold - Eval)
        inclvl = self._itcount > 0 and diff < self._deltaE
        self._Eold = Eval
        self._ccount = self._ccount + 1 if inclvl else max(0, self._ccount - 1)
        if self._iteration_limit is not None and self._itcount >= self._iteration_limit:
            return CONVERGED
        if self._ccount >= self._convergence_level:
            return CONVERGED
        return CONTINUE


def conjugate_gradient_np(energy, controller, preconditioner=None, nreset=20, history=None, res_history=None):
    """``ConjugateGradient.__call__`` (``conjugate_gradient.py:48-126``).

    Returns ``(energy, status, n_iterations)``; ``history`` (a list) receives the
    energy value after every step, ``res_history`` the gradient norm.
    """
    status = controller.start(energy)
    if status != CONTINUE:
        return energy, status, 0
    r = energy.gradient
    d = r if preconditioner is None else preconditioner(r)
    previous_gamma = _vdot(r, d)
    if np.isnan(previous_gamma):
        return energy, ERROR, 0
    if previous_gamma == 0:
        return energy, CONVERGED, 0
    ii = nit = 0
    while True:
        q = energy.apply_metric(d)
        curv = _vdot(d, q)
        if np.isnan(curv) or curv == 0.0:
            return energy, ERROR, nit
        alpha = previous_gamma / curv
        if alpha < 0:
            return energy, ERROR, nit
        ii += 1
        nit += 1
        if ii < nreset:
            r = r - q * alpha
            energy = energy.at_with_grad(energy.position - alpha * d, r)
        else:
            energy = energy.at(energy.position - alpha * d)
            r = energy.gradient
            ii = 0
        s = r if preconditioner is None else preconditioner(r)
        gamma = _vdot(r, s)
        if history is not None:
            history.append(energy.value)
        if res_history is not None:
            res_history.append(energy.gradient_norm)
        if np.isnan(gamma) or gamma < 0:
            return energy, ERROR, nit
        if gamma == 0:
            return energy, CONVERGED, nit
        status = controller.check(energy)
        if status != CONTINUE:
            return energy, status, nit
        d = d * max(0, gamma / previous_gamma) + s
        previous_gamma = gamma


# -------------------------------------------------------------------- re stack
CGResults = namedtuple("CGResults", "x nit nfev info success")


def _defaults(j, miniter, maxiter, absdelta, resnorm, norm_ord, tol, atol):
    """``re/conjugate_gradient.py:93-107``."""
    norm_ord = 2 if norm_ord is None else norm_ord
    maxiter_fallback = 20 * j.size
    miniter = min(6, maxiter if maxiter is not None else maxiter_fallback) if miniter is None else miniter
    maxiter = max(min(200, maxiter_fallback), miniter) if maxiter is None else maxiter
    if absdelta is None and resnorm is None:
        resnorm = max(tol * float(np.linalg.norm(np.ravel(j), ord=norm_ord)), atol)
    fi = np.finfo(j.dtype if np.issubdtype(j.dtype, np.floating) else np.float64)
    return norm_ord, miniter, maxiter, resnorm, 6.0 * fi.eps, 6.0 * fi.tiny


def cg_re(mat, j, x0=None, *, absdelta=None, resnorm=None, norm_ord=None, tol=1e-5, atol=0.0,
          miniter=None, maxiter=None, _raise_nonposdef=True, history=None, res_history=None):
    """``_cg`` (``re/conjugate_gradient.py:77-216``), Python-loop variant.  ``history`` / ``res_history``
    (lists) receive the energy / the residual norm after every step (test instrumentation)."""
    norm_ord, miniter, maxiter, resnorm, eps, tiny = _defaults(
        j, miniter, maxiter, absdelta, resnorm, norm_ord, tol, atol)
    if x0 is None:
        pos, r, energy, nfev = np.zeros_like(j), -j, 0.0, 0
    else:
        pos = x0
        r = mat(pos) - j
        energy, nfev = _vdot((r - j) / 2, pos), 1
    d = r
    previous_gamma = _vdot(r, r)
    info, i = -1, 0
    if previous_gamma == 0:
        return CGResults(pos, 0, nfev, 0, True)
    for i in range(1, maxiter + 1):
        q = mat(d)
        nfev += 1
        curv = _vdot(d, q)
        if curv == 0.0:
            if _raise_nonposdef:
                raise ValueError("CG: zero curvature")
            info = 0
            break
        elif curv < 0.0:
            if _raise_nonposdef:
                raise ValueError("CG: negative curvature")
            if i > 1:
                info = 0
                break
            pos = previous_gamma / (-curv) * j
            info = 0
            break
        alpha = previous_gamma / curv
        pos = pos - alpha * d
        if i % N_RESET == 0:
            r = mat(pos) - j
            nfev += 1
        else:
            r = r - q * alpha
        gamma = _vdot(r, r)
        if res_history is not None:
            res_history.append(float(np.linalg.norm(np.ravel(r), ord=norm_ord)))
        if gamma >= 0.0 and gamma <= tiny:
            info = 0
            break
        if resnorm is not None:
            norm = float(np.linalg.norm(np.ravel(r), ord=norm_ord))
            if norm < resnorm and i >= miniter:
                info = 0
                break
        new_energy = _vdot((r - j) / 2, pos)
        energy_diff = energy - new_energy
        if history is not None:
            history.append(new_energy)
        if energy_diff < -eps * abs(new_energy):
            if _raise_nonposdef:
                raise ValueError("CG: WARNING: energy increased")
            info = i
            break
        if absdelta is not None and energy_diff < absdelta and i >= miniter:
            info = 0
            break
        energy = new_energy
        d = d * max(0, gamma / previous_gamma) + r
        previous_gamma = gamma
    info = i if info == -1 else info
    return CGResults(pos, i, nfev, info, info == 0)


def static_cg_re(mat, j, x0=None, *, absdelta=None, resnorm=None, norm_ord=None, tol=1e-5,
                 atol=0.0, miniter=None, maxiter=None, history=None):
    """``_static_cg`` (``re/conjugate_gradient.py:219-389``): the ``lax.while_loop``
    body ``cg_single_step`` (``:266-338``) executed as a host loop with identical
    ``info`` bookkeeping (-2 running, -1 error, 0 converged, i>0 maxiter)."""
    norm_ord, miniter, maxiter, resnorm, eps, tiny = _defaults(
        j, miniter, maxiter, absdelta, resnorm, norm_ord, tol, atol)
    if x0 is None:
        pos, r, nfev, energy = np.zeros_like(j), -j, 0, 0.0
    else:
        pos = x0
        r = mat(pos) - j
        nfev, energy = 1, _vdot((r - j) / 2, pos)
    d = r
    gamma = _vdot(r, r)
    info, i = (0 if gamma == 0.0 else -2), 0
    while info < -1:
        previous_gamma, previous_energy = gamma, energy
        i += 1
        q = mat(d)
        curv = _vdot(d, q)
        if curv == 0.0:
            info = -1
        with np.errstate(divide="ignore", invalid="ignore"):
            alpha = np.float64(previous_gamma) / np.float64(curv)
        if alpha < 0.0:
            info = -1
        pos = pos - alpha * d
        r = (mat(pos) - j) if i % N_RESET == 0 else (r - q * alpha)
        gamma = _vdot(r, r)
        if gamma >= 0.0 and gamma <= tiny and info != -1:
            info = 0
        if resnorm is not None:
            norm = float(np.linalg.norm(np.ravel(r), ord=norm_ord))
            if norm < resnorm and i >= miniter and info != -1:
                info = 0
        energy = _vdot((r - j) / 2, pos)
        energy_diff = previous_energy - energy
        if history is not None:
            history.append(energy)
        if energy_diff < -eps * abs(energy):
            info = -1
        if absdelta is not None and energy_diff < absdelta and i >= miniter and info != -1:
            info = 0
        if i >= maxiter and info != -1:
            info = i if info == -2 else info
            # ``jnp.where((i >= maxiter) & (info != -1), i, info)`` (:313) overwrites even a 0
            info = i
        with np.errstate(divide="ignore", invalid="ignore"):
            d = d * max(0.0, np.float64(gamma) / np.float64(previous_gamma)) + r
    nfev += i + i // N_RESET
    return CGResults(pos, i, nfev, info, info == 0)
