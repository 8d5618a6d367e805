"""Conjugate-gradient solvers with fused device kernels.

Two interfaces, matching the two reference stacks:

* :class:`ConjugateGradient` + :class:`QuadraticEnergy` + iteration controllers -- numpy stack
  (``/root/reference/src/minimization/conjugate_gradient.py:48-126``,
  ``src/minimization/quadratic_energy.py:25-78``,
  ``src/minimization/iteration_controllers.py:148-221,356-423``);
* :func:`cg` / :func:`static_cg` -- NIFTy.re (``src/re/conjugate_gradient.py:77-216,219-389``).

Vectors are device tensors.  Per iteration: one metric apply (the curvature ``d.q`` is fused
into its last store when the operator offers ``apply_with_curvature``), ONE fused kernel for
``x -= a d, r -= a q`` together with the sums ``r.r, x.r, x.b`` (fp64, deterministic), one
kernel for ``d = b d + r``.  The stopping logic is evaluated on the host in fp64 from those
three sums, line by line as in the reference, so iteration counts and status/info codes match.
"""

import datetime
from collections import namedtuple

import numpy as np
import torch

from . import _lib
from .transforms import _TORCH_REAL, _ptr, _stream, vdot_device

CONVERGED, CONTINUE, ERROR = 0, 1, 2   # IterationController.{CONVERGED,CONTINUE,ERROR}
N_RESET = 20                           # re/conjugate_gradient.py:18


# ---------------------------------------------------------------- fused vector kernels
def _dot(a, b, comm=None):
    """sum(a*b) as a Python float; all-reduced over the ranks of a slab-decomposed vector."""
    t = vdot_device(a, b)
    if comm is not None:
        comm.allreduce_(t)
    return float(t.cpu())


class _Vec:
    def __init__(self, like, comm=None):
        self.comm = comm
        self.lib = _lib.get()
        if like.dtype not in _TORCH_REAL:
            raise TypeError("CG vectors must be float32 or float64 tensors")
        self.code, self.n = _TORCH_REAL[like.dtype], like.numel()
        self.sums = torch.zeros(3, dtype=torch.float64, device=like.device)

    def update(self, mode, alpha, d, q, b, x, r):
        """mode 0: x -= alpha d, r -= alpha q; mode 1: r = q - b.  Returns (r.r, x.r, x.b)."""
        lib = self.lib
        rc = lib.nb2_cg_update(self.code, self.n, mode, float(alpha), _ptr(lib, d), _ptr(lib, q), _ptr(lib, b),
                               _ptr(lib, x), _ptr(lib, r), _ptr(lib, self.sums), _stream(x))
        lib.check(rc, "nb2_cg_update")
        if self.comm is not None:
            self.comm.allreduce_(self.sums)      # one all-reduce of three fp64 scalars
        s = self.sums.cpu()
        return float(s[0]), float(s[1]), float(s[2])

    def axpby(self, a, x, b, y):
        """y = a x + b y"""
        lib = self.lib
        lib.check(lib.nb2_axpby(self.code, self.n, float(a), _ptr(lib, x), float(b), _ptr(lib, y), _stream(y)),
                  "nb2_axpby")
        return y


def _apply_with_curv(A, d, q, pending=None, vec=None):
    """``q = A d`` and ``d.q``.  ``pending = (s, beta)``: the deferred direction update ``d <- s + beta d``
    of the previous iteration, done inside the first pass of ``A`` when it can carry it
    (``A.fused_direction_update``), else by the axpby kernel right here."""
    fused = getattr(A, "fused_direction_update", False) and hasattr(A, "apply_with_curvature")
    if pending is not None and not fused:
        vec.axpby(1.0, pending[0], pending[1], d)
        pending = None
    if hasattr(A, "apply_with_curvature"):
        q, curv = A.apply_with_curvature(d, out=q, pre=pending) if pending is not None else \
            A.apply_with_curvature(d, out=q)
        return q, float(curv.cpu())
    res = A(d)
    if res.data_ptr() != q.data_ptr():
        q.copy_(res)
    return q, _dot(d, q, getattr(A, "comm", None))


def _apply(A, x, out):
    try:
        res = A(x, out=out)
    except TypeError:
        res = A(x)
    if res.data_ptr() != out.data_ptr():
        out.copy_(res)
    return out


# ================================================================ numpy-stack interface
class QuadraticEnergy:
    """``E(x) = 0.5 x^T A x - b^T x`` (quadratic_energy.py:25-78) on device tensors."""

    def __init__(self, position, A, b, _grad=None, _value=None, _gradnorm=None):
        self.position, self._A, self._b = position, A, b
        if _grad is not None:
            self.gradient = _grad
            Ax = None
        else:
            Ax = A(position)
            self.gradient = Ax if b is None else Ax - b
        if _value is None:
            # 0.5 x.(A x) - b.x with A x = grad + b (quadratic_energy.py:32-37); no N-sized temporary
            if Ax is not None:
                _value = 0.5 * float(vdot_device(position, Ax).cpu())
            else:
                _value = 0.5 * float(vdot_device(position, _grad).cpu())
                if b is not None:
                    _value += 0.5 * float(vdot_device(position, b).cpu())
            if b is not None:
                _value -= float(vdot_device(b, position).cpu())
        self.value = _value
        self._gradnorm = _gradnorm

    def at(self, position):
        return QuadraticEnergy(position, self._A, self._b)

    def at_with_grad(self, position, grad):
        return QuadraticEnergy(position, self._A, self._b, grad)

    @property
    def gradient_norm(self):
        if self._gradnorm is None:
            self._gradnorm = float(np.sqrt(float(vdot_device(self.gradient, self.gradient).cpu())))
        return self._gradnorm

    def apply_metric(self, x):
        return self._A(x)

    @property
    def metric(self):
        return self._A


class _EnergyView:
    """What an iteration controller looks at (value, gradient_norm), from the fused sums."""

    __slots__ = ("value", "gradient_norm")

    def __init__(self, value, gradient_norm):
        self.value, self.gradient_norm = value, gradient_norm


class IterationController:
    CONVERGED, CONTINUE, ERROR = CONVERGED, CONTINUE, ERROR


class GradientNormController(IterationController):
    """``iteration_controllers.py:148-221``."""

    def __init__(self, tol_abs_gradnorm=None, tol_rel_gradnorm=None, convergence_level=1, iteration_limit=None,
                 name=None):
        self._tol_abs_gradnorm, self._tol_rel_gradnorm = tol_abs_gradnorm, tol_rel_gradnorm
        self._convergence_level, self._iteration_limit, self._name = convergence_level, iteration_limit, name

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


class AbsDeltaEnergyController(IterationController):
    """``iteration_controllers.py:356-423``."""

    def __init__(self, deltaE, convergence_level=1, iteration_limit=None, name=None):
        self._deltaE, self._convergence_level = deltaE, convergence_level
        self._iteration_limit, self._name = iteration_limit, name

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


class ConjugateGradient:
    """``ConjugateGradient`` (conjugate_gradient.py:27-126): same constructor, same call,
    same return ``(energy, status)``; ``nit`` of the last call is kept in ``self.nit``."""

    def __init__(self, controller, nreset=20):
        self._controller, self._nreset = controller, nreset
        self.nit = 0

    def __call__(self, energy, preconditioner=None):
        controller = self._controller
        self.nit = 0
        status = controller.start(energy)
        if status != CONTINUE:
            return energy, status
        A, b = energy._A, energy._b
        if b is None:
            b = torch.zeros_like(energy.position)
        vec = _Vec(energy.position)
        x = energy.position.clone()
        r = energy.gradient.clone()
        d = r.clone() if preconditioner is None else preconditioner(r).clone()
        q = torch.empty_like(x)
        previous_gamma = float(vdot_device(r, d).cpu())
        if np.isnan(previous_gamma):
            return energy, ERROR
        if previous_gamma == 0:
            return energy, CONVERGED

        def mk_energy(view):
            # materialised on return only; x and r are buffers owned by this call
            return QuadraticEnergy(x, A, energy._b, _grad=r, _value=view.value, _gradnorm=view.gradient_norm)

        ii = 0
        pending = None    # deferred direction update (s, beta): done with the next A d
        while True:
            q, curv = _apply_with_curv(A, d, q, pending, vec)         # :84-85 (+ the deferred :124)
            pending = None
            if np.isnan(curv) or curv == 0.0:                          # :86-91
                return energy, ERROR
            alpha = previous_gamma / curv
            if alpha < 0:                                              # :93-96
                return energy, ERROR
            ii += 1
            self.nit += 1
            if ii < self._nreset:                                      # :99-101
                rr, xr, xb = vec.update(0, alpha, d, q, b, x, r)
            else:                                                      # :102-105 exact residual
                vec.axpby(-alpha, d, 1.0, x)
                q = _apply(A, x, q)
                rr, xr, xb = vec.update(1, 0.0, d, q, b, x, r)
                ii = 0
            value = 0.5 * (xr + xb) - xb                               # quadratic_energy.py:35-37
            view = _EnergyView(value, float(np.sqrt(rr)))              # what the controller looks at
            if preconditioner is None:
                s, gamma = r, rr
            else:
                s = preconditioner(r)
                gamma = float(vdot_device(r, s).cpu())
            if np.isnan(gamma) or gamma < 0:                           # :110-115
                return mk_energy(view), ERROR
            if gamma == 0:
                return mk_energy(view), CONVERGED
            status = controller.check(view)                            # :120
            if status != CONTINUE:
                return mk_energy(view), status
            pending = (s, max(0.0, gamma / previous_gamma))            # :124, applied with the next A d
            previous_gamma = gamma


# ================================================================ NIFTy.re interface
CGResults = namedtuple("CGResults", "x nit nfev info success")


def _defaults(j, miniter, maxiter, absdelta, resnorm, norm_ord, tol, atol, comm=None):
    """re/conjugate_gradient.py:93-107."""
    norm_ord = 2 if norm_ord is None else norm_ord
    maxiter_fallback = 20 * j.numel() * (comm.size if comm is not None else 1)
    miniter = min(6, maxiter if maxiter is not None else maxiter_fallback) if miniter is None else miniter
    maxiter = max(min(200, maxiter_fallback), miniter) if maxiter is None else maxiter
    if absdelta is None and resnorm is None:
        resnorm = max(tol * _norm(j, None, norm_ord, comm), atol)
    fi = np.finfo(np.float32 if j.dtype == torch.float32 else np.float64)
    return norm_ord, miniter, maxiter, resnorm, 6.0 * fi.eps, 6.0 * fi.tiny


def _norm(r, rr, norm_ord, comm=None):
    """``jft.norm(r, ord)`` (re/tree_math/vector_math.py:176-192); ``rr`` = r.r when already known.  On a
    slab-decomposed vector every rank gets the GLOBAL norm (all ranks must take the same stopping decision)."""
    if norm_ord == 2:
        return float(np.sqrt(rr if rr is not None else _dot(r, r, comm)))
    if comm is None:
        return float(torch.linalg.vector_norm(r.reshape(-1).double(), ord=norm_ord).cpu())
    a = r.reshape(-1).double().abs()
    if norm_ord == np.inf:
        t = a.max().reshape(1)
        comm.allreduce_max_(t)
        return float(t.cpu())
    if norm_ord == -np.inf:
        t = (-a.min()).reshape(1)
        comm.allreduce_max_(t)
        return -float(t.cpu())
    t = ((a != 0).sum().double() if norm_ord == 0 else (a ** norm_ord).sum()).reshape(1)
    comm.allreduce_(t)
    return float(t.cpu()) if norm_ord == 0 else float(t.cpu()) ** (1.0 / norm_ord)


def _timed_out(time_threshold, comm=None):
    """``datetime.now() > time_threshold`` (re/conjugate_gradient.py:176); a float is read as a
    ``time.time()`` stamp.  Slab-decomposed solves take rank 0's decision (one broadcast)."""
    if time_threshold is None:
        return False
    if isinstance(time_threshold, datetime.datetime):
        out = datetime.datetime.now() > time_threshold
    else:
        out = datetime.datetime.now().timestamp() > float(time_threshold)
    if comm is not None:
        out = comm.broadcast_flag(out)
    return out


def _cg_setup(mat, j, x0):
    comm = getattr(mat, "comm", None)
    vec = _Vec(j, comm)
    q = torch.empty_like(j)
    if x0 is None:
        pos = torch.zeros_like(j)
        r = -j
        energy, nfev = 0.0, 0
        gamma = _dot(r, r, comm)
    else:
        pos = x0.clone()
        q = _apply(mat, pos, q)
        r = torch.empty_like(j)
        rr, xr, xb = vec.update(1, 0.0, pos, q, j, pos, r)       # r = mat(x0) - j
        energy, nfev, gamma = 0.5 * (xr - xb), 1, rr             # vdot((r - j)/2, pos)  :118-121
    d = r.clone()
    return vec, pos, r, d, q, energy, nfev, gamma


def cg(mat, j, x0=None, *, absdelta=None, resnorm=None, norm_ord=None, tol=1e-5, atol=0.0, miniter=None,
       maxiter=None, name=None, time_threshold=None, _raise_nonposdef=True, history=None):
    """``nifty8.re.conjugate_gradient.cg`` / ``_cg`` (re/conjugate_gradient.py:29-42,77-216).

    Returns ``(x, CGResults)`` like the public ``cg`` (``:29-42``: ``return cg_res.x, cg_res.info``
    is the reference's public signature; the full record is returned here for inspection)."""
    res = _cg(mat, j, x0, absdelta=absdelta, resnorm=resnorm, norm_ord=norm_ord, tol=tol, atol=atol,
              miniter=miniter, maxiter=maxiter, name=name, time_threshold=time_threshold,
              _raise_nonposdef=_raise_nonposdef, history=history)
    return res.x, res.info


def _cg(mat, j, x0=None, *, absdelta=None, resnorm=None, norm_ord=None, tol=1e-5, atol=0.0, miniter=None,
        maxiter=None, name=None, time_threshold=None, _raise_nonposdef=True, history=None, res_history=None):
    """``_cg`` (re/conjugate_gradient.py:77-216).  ``history`` / ``res_history``: optional lists that receive
    the energy / the residual norm after every iteration (instrumentation, not in the reference)."""
    norm_ord, miniter, maxiter, resnorm, eps, tiny = _defaults(j, miniter, maxiter, absdelta, resnorm, norm_ord,
                                                               tol, atol, getattr(mat, "comm", None))
    vec, pos, r, d, q, energy, nfev, previous_gamma = _cg_setup(mat, j, x0)
    info, i = -1, 0
    pending = None    # deferred direction update (r, beta): done with the next mat(d)
    if previous_gamma == 0:
        return CGResults(pos, 0, nfev, 0, True)
    for i in range(1, maxiter + 1):
        q, curv = _apply_with_curv(mat, d, q, pending, vec)            # :147-150 (+ the deferred :205)
        pending = None
        nfev += 1
        if curv == 0.0:                                                # :151-156
            if _raise_nonposdef:
                raise ValueError("zero curvature in conjugate gradient")
            info = 0
            break
        elif curv < 0.0:                                               # :157-167
            if _raise_nonposdef:
                raise ValueError("negative curvature in conjugate gradient")
            if i > 1:
                info = 0
                break
            pos = previous_gamma / (-curv) * j
            info = 0
            break
        alpha = previous_gamma / curv
        if i % N_RESET == 0:                                           # :170-172
            vec.axpby(-alpha, d, 1.0, pos)
            q = _apply(mat, pos, q)
            nfev += 1
            gamma, xr, xb = vec.update(1, 0.0, d, q, j, pos, r)
        else:                                                          # :169,174
            gamma, xr, xb = vec.update(0, alpha, d, q, j, pos, r)
        if _timed_out(time_threshold, vec.comm):                       # :176-178
            info = i
            break
        if res_history is not None:
            res_history.append(_norm(r, gamma, norm_ord, vec.comm))
        if gamma >= 0.0 and gamma <= tiny:                             # :179-183
            info = 0
            break
        if resnorm is not None:                                        # :184-188
            if _norm(r, gamma, norm_ord, vec.comm) < resnorm and i >= miniter:
                info = 0
                break
        new_energy = 0.5 * (xr - xb)                                   # :191
        energy_diff = energy - new_energy
        if history is not None:
            history.append(new_energy)
        if energy_diff < -eps * abs(new_energy):                       # :192-200
            if _raise_nonposdef:
                raise ValueError("WARNING: energy increased in conjugate gradient")
            info = i
            break
        if absdelta is not None and energy_diff < absdelta and i >= miniter:   # :201-203
            info = 0
            break
        energy = new_energy
        pending = (r, max(0.0, gamma / previous_gamma))                # :205, applied with the next mat(d)
        previous_gamma = gamma
    info = i if info == -1 else info
    return CGResults(pos, i, nfev, info, info == 0)


def static_cg(mat, j, x0=None, *, absdelta=None, resnorm=None, norm_ord=None, tol=1e-5, atol=0.0, miniter=None,
              maxiter=None, name=None, history=None, **kwargs):
    """``static_cg`` (re/conjugate_gradient.py:45-51): returns ``(x, info)``."""
    res = _static_cg(mat, j, x0, absdelta=absdelta, resnorm=resnorm, norm_ord=norm_ord, tol=tol, atol=atol,
                     miniter=miniter, maxiter=maxiter, name=name, history=history)
    return res.x, res.info


def _static_cg(mat, j, x0=None, *, absdelta=None, resnorm=None, norm_ord=None, tol=1e-5, atol=0.0, miniter=None,
               maxiter=None, name=None, history=None, device_resident=None, poll=4, **kwargs):
    """``_static_cg`` (re/conjugate_gradient.py:219-389): the body of ``cg_single_step``
    (``:266-338``) with the same ``info`` bookkeeping (-2 running, -1 error, 0 converged,
    i > 0 iteration limit).

    Device-resident by default, like the reference's ``lax.while_loop`` (:382): scalars, stopping tests and
    alpha / beta stay on the device (:class:`_DeviceCG`), the host polls ``info`` every ``poll`` iterations.
    The host-driven loop below (one read-back of 4 scalars per iteration) is used when per-iteration values
    are wanted on the host (``history``, ``name`` = logging, as the reference's host callback :317-327), for
    ``norm_ord != 2`` and for operators without ``apply_with_curvature``."""
    norm_ord, miniter, maxiter, resnorm, eps, tiny = _defaults(j, miniter, maxiter, absdelta, resnorm, norm_ord,
                                                               tol, atol, getattr(mat, "comm", None))
    if device_resident is None:
        device_resident = (history is None and name is None and norm_ord == 2
                           and hasattr(mat, "apply_with_curvature") and getattr(mat, "device_resident_cg", False))
    if device_resident:
        return _DeviceCG(mat, j, x0, miniter, maxiter, resnorm, absdelta, eps, tiny, poll=poll).run()
    vec, pos, r, d, q, energy, nfev, gamma = _cg_setup(mat, j, x0)
    info, i = (0 if gamma == 0.0 else -2), 0
    pending = None    # deferred direction update (r, beta): done with the next mat(d)
    while info < -1:
        previous_gamma, previous_energy = gamma, energy
        i += 1
        q, curv = _apply_with_curv(mat, d, q, pending, vec)            # :270-271 (+ the deferred :315)
        pending = None
        if curv == 0.0:
            info = -1
        with np.errstate(divide="ignore", invalid="ignore"):
            alpha = float(np.float64(previous_gamma) / np.float64(curv))
        if alpha < 0.0:                                                # :276-279
            info = -1
        if i % N_RESET == 0:                                           # :281-290
            vec.axpby(-alpha, d, 1.0, pos)
            q = _apply(mat, pos, q)
            gamma, xr, xb = vec.update(1, 0.0, d, q, j, pos, r)
        else:
            gamma, xr, xb = vec.update(0, alpha, d, q, j, pos, r)
        if gamma >= 0.0 and gamma <= tiny and info != -1:              # :292-294
            info = 0
        if resnorm is not None:                                        # :295-301
            if _norm(r, gamma, norm_ord, vec.comm) < resnorm and i >= miniter and info != -1:
                info = 0
        energy = 0.5 * (xr - xb)                                       # :303
        energy_diff = previous_energy - energy
        if history is not None:
            history.append(energy)
        if energy_diff < -eps * abs(energy):                           # :305-307
            info = -1
        if absdelta is not None and energy_diff < absdelta and i >= miniter and info != -1:   # :308-312
            info = 0
        if i >= maxiter and info != -1:                                # :313
            info = i
        with np.errstate(divide="ignore", invalid="ignore"):
            beta = max(0.0, float(np.float64(gamma) / np.float64(previous_gamma)))
        pending = (r, beta)                                            # :315, applied with the next mat(d)
    nfev += i + i // N_RESET
    return CGResults(pos, i, nfev, info, info == 0)


# ---------------------------------------------------------------- device-resident static_cg
class _DeviceCG:
    """The reference's ``lax.while_loop`` CG (re/conjugate_gradient.py:219-389) with the loop state on the
    device: the scalars of the recurrence live in a device array (``nb2_cg_decide``), alpha / beta are read
    from it by the update kernels and the fused first pass, and the stopping tests run in a one-thread kernel.
    The host only enqueues iterations and polls ``info`` every ``poll`` iterations; iterations enqueued
    past the stopping decision are no-ops (launch gate).  On a single GPU the regular iteration is captured
    once as a CUDA graph and replayed."""

    def __init__(self, mat, j, x0, miniter, maxiter, resnorm, absdelta, eps, tiny, poll=4, use_graph=None):
        self.mat, self.j = mat, j
        self.lib = lib = _lib.get()
        self.comm = getattr(mat, "comm", None)
        self.vec, self.pos, self.r, self.d, self.q, energy, self.nfev0, gamma = _cg_setup(mat, j, x0)
        self.code, self.n = self.vec.code, self.vec.n
        dev = j.device
        st = np.zeros(_lib.CG_STATE_SIZE)
        st[_lib.CG_GAMMA], st[_lib.CG_ENERGY] = gamma, energy
        st[_lib.CG_INFO] = 0 if gamma == 0.0 else -2                   # :363-366
        st[_lib.CG_MINITER], st[_lib.CG_MAXITER] = miniter, maxiter
        st[_lib.CG_RESNORM] = -1.0 if resnorm is None else resnorm
        st[_lib.CG_ABSDELTA] = np.nan if absdelta is None else absdelta
        st[_lib.CG_EPS], st[_lib.CG_TINY] = eps, tiny
        self.state = torch.from_numpy(st).to(dev)
        self.gate = torch.zeros(1, dtype=torch.int32, device=dev)
        if gamma == 0.0:
            self.gate.fill_(1)
        self.curv = self.state[_lib.CG_CURV:_lib.CG_CURV + 1]
        self.beta = self.state[_lib.CG_BETA:_lib.CG_BETA + 1]
        self.alpha = self.state[_lib.CG_ALPHA:_lib.CG_ALPHA + 1]
        self.sums = self.state[_lib.CG_SUMS:_lib.CG_SUMS + 3]
        self.fused = getattr(mat, "fused_direction_update", False)
        # single process: the scalar step runs inside the launch of the reduction that feeds it (nb2_set_cg_fusion)
        self.fuse_decide = self.comm is None and bool(getattr(mat, "fuses_cg_decide", False))
        self.poll = max(1, int(poll))
        if use_graph is None:
            use_graph = dev.type == "cuda" and self.comm is None
        self.use_graph, self.graph = use_graph, None
        self.miniter, self.maxiter = int(miniter), int(maxiter)
        self.i = 0                       # iterations enqueued so far
        self.graph_kernels = 0           # kernels inside the captured iteration
        self.replayed_kernels = 0        # kernels launched through graph replays (not seen by nb2_launch_count)

    def _decide(self, stage):
        lib = self.lib
        lib.check(lib.nb2_cg_decide(_ptr(lib, self.state), _ptr(lib, self.gate), stage, _stream(self.state)),
                  "nb2_cg_decide")

    def _iteration(self, i):
        """Enqueue iteration ``i`` (1-based); nothing is read back."""
        lib, mat, j = self.lib, self.mat, self.j
        pending = i > 1                                   # d <- r + beta d of the previous iteration (:315)
        if pending and not self.fused:
            lib.check(lib.nb2_axpby_dev(self.code, self.n, 1.0, None, _ptr(lib, self.r), 1.0, _ptr(lib, self.beta),
                                        _ptr(lib, self.d), _stream(self.d)), "nb2_axpby_dev")
        pre = (self.r, self.beta) if (pending and self.fused) else None
        if pre is not None:
            mat.apply_with_curvature(self.d, out=self.q, pre=pre, curv_out=self.curv)     # :270-271
        else:
            mat.apply_with_curvature(self.d, out=self.q, curv_out=self.curv)
        if not self.fuse_decide:
            self._decide(0)                               # alpha = gamma / curv (:272-276)
        if i % N_RESET == 0:                              # :277-290 exact residual
            lib.check(lib.nb2_axpby_dev(self.code, self.n, -1.0, _ptr(lib, self.alpha), _ptr(lib, self.d), 1.0, None,
                                        _ptr(lib, self.pos), _stream(self.pos)), "nb2_axpby_dev")
            _apply(mat, self.pos, self.q)
            mode = 1
        else:
            mode = 0
        lib.check(lib.nb2_cg_update_dev(self.code, self.n, mode, _ptr(lib, self.state), _ptr(lib, self.d),
                                        _ptr(lib, self.q), _ptr(lib, j), _ptr(lib, self.pos), _ptr(lib, self.r),
                                        _stream(self.pos)), "nb2_cg_update_dev")
        if self.comm is not None:
            self.comm.allreduce_(self.sums)
        if not self.fuse_decide:
            self._decide(1)                               # :291-315 + continue_condition

    def _regular(self, i):
        """Iterations with the pending update and no reset are identical launches: replay a graph."""
        if not self.use_graph:
            return self._iteration(i)
        if self.graph is None:
            n0 = self.lib.nb2_launch_count()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self._iteration(i)
            self.graph, self.graph_kernels = g, int(self.lib.nb2_launch_count() - n0)
        self.graph.replay()
        self.replayed_kernels += self.graph_kernels

    def enqueue(self, n):
        """Enqueue the next ``n`` iterations without reading anything back."""
        for _ in range(n):
            self.i += 1
            if self.i > 1 and self.i % N_RESET != 0:
                self._regular(self.i)
            else:
                self._iteration(self.i)

    def hooks(self, on):
        """Launch gate (and, single process, the fused scalar step) of the calling thread on / off."""
        lib = self.lib
        if on:
            lib.check(lib.nb2_set_launch_gate(_ptr(lib, self.gate)), "nb2_set_launch_gate")
            if self.fuse_decide:
                lib.check(lib.nb2_set_cg_fusion(_ptr(lib, self.state), _ptr(lib, self.gate)), "nb2_set_cg_fusion")
        else:
            lib.nb2_set_launch_gate(None)
            lib.nb2_set_cg_fusion(None, None)

    def run(self):
        lib = self.lib
        info = int(self.state[_lib.CG_INFO].item())
        self.hooks(True)
        try:
            while info < -1 and self.i < self.maxiter:
                self.enqueue(1)
                # nothing but an error can stop the loop before miniter: poll from there on
                if self.i >= min(self.miniter, self.maxiter) and (self.i % self.poll == 0 or self.i == self.maxiter):
                    info = int(self.state[_lib.CG_INFO].item())
        finally:
            self.hooks(False)
        st = self.state.cpu().numpy()
        nit, info = int(st[_lib.CG_ITER]), int(st[_lib.CG_INFO])
        nfev = self.nfev0 + nit + nit // N_RESET
        return CGResults(self.pos, nit, nfev, info, info == 0)


class CGLoop:
    """The CG recurrence with the stopping tests taken out: ``run(n)`` executes exactly ``n``
    iterations with the same primitives as the solvers (metric apply with fused curvature,
    fused update + sums, direction update; exact residual every ``N_RESET`` iterations).
    Used by ``bench.py`` / profiling, where a fixed amount of work has to be timed.

    ``device_resident`` (default when the operator allows it): the loop of :func:`static_cg` -- scalars on the
    device, nothing read back, the regular iteration replayed as a CUDA graph on a single GPU; ``run`` only
    enqueues.  Otherwise the host-driven loop of :func:`cg` (4 scalars read back per iteration)."""

    def __init__(self, mat, j, x0=None, device_resident=None):
        self.mat, self.j = mat, j
        if device_resident is None:
            device_resident = hasattr(mat, "apply_with_curvature") and getattr(mat, "device_resident_cg", False)
        self.dev = None
        if device_resident:
            # every criterion off: only an error (non-positive curvature, energy increase) could stop it
            self.dev = _DeviceCG(mat, j, x0, 2 ** 40, 2 ** 40, None, None, 0.0, 0.0)
            self.dev.state[_lib.CG_EPS] = float("inf")
            self.vec, self.pos, self.r, self.d, self.q = self.dev.vec, self.dev.pos, self.dev.r, self.dev.d, self.dev.q
            return
        (self.vec, self.pos, self.r, self.d, self.q, self.energy, self.nfev, self.gamma) = _cg_setup(mat, j, x0)
        self.i = 0
        self.pending = None    # deferred direction update (r, beta), applied with the next metric apply

    @property
    def replayed_kernels(self):
        return self.dev.replayed_kernels if self.dev is not None else 0

    def run(self, n):
        if self.dev is not None:
            self.dev.hooks(True)
            try:
                self.dev.enqueue(n)
            finally:
                self.dev.hooks(False)
            return self
        mat, vec, j = self.mat, self.vec, self.j
        for _ in range(n):
            self.i += 1
            self.q, curv = _apply_with_curv(mat, self.d, self.q, self.pending, vec)
            self.pending = None
            self.nfev += 1
            alpha = self.gamma / curv
            if self.i % N_RESET == 0:
                vec.axpby(-alpha, self.d, 1.0, self.pos)
                self.q = _apply(mat, self.pos, self.q)
                self.nfev += 1
                gamma, xr, xb = vec.update(1, 0.0, self.d, self.q, j, self.pos, self.r)
            else:
                gamma, xr, xb = vec.update(0, alpha, self.d, self.q, j, self.pos, self.r)
            self.energy = 0.5 * (xr - xb)
            self.pending = (self.r, max(0.0, gamma / self.gamma))
            self.gamma = gamma
        return self
