"""Oracle: ``_newton_cg`` (``/root/reference/src/re/optimize.py:102-240``) restated on NumPy vectors
(test infrastructure; jax is absent here).  Uses the oracle's ``cg_re`` for the inner solves."""

from collections import namedtuple

import numpy as np

from . import cg as _cg

OptimizeResults = namedtuple("OptimizeResults", "x success status fun jac nit nfev njev nhev")


def newton_cg(fun_and_grad, x0, hessp, *, miniter=None, maxiter=None, energy_reduction_factor=0.1, old_fval=None,
              absdelta=None, norm_ord=None, xtol=1e-5, cg_kwargs=None, custom_gradnorm=None):
    norm_ord = 1 if norm_ord is None else norm_ord
    miniter = 0 if miniter is None else miniter
    maxiter = 200 if maxiter is None else maxiter
    xtol = xtol * x0.size
    pos = x0
    cg_kwargs = {} if cg_kwargs is None else cg_kwargs
    gradnorm = (lambda v: float(np.linalg.norm(np.ravel(v), ord=norm_ord))) if custom_gradnorm is None else custom_gradnorm
    energy, g = fun_and_grad(pos)
    nfev, njev, nhev = 1, 1, 0
    if np.isnan(energy):
        raise ValueError("energy is Nan")
    status, i = -1, 0
    for i in range(1, maxiter + 1):
        if old_fval and energy_reduction_factor:
            cg_absdelta = energy_reduction_factor * (old_fval - energy)
        else:
            cg_absdelta = None if absdelta is None else absdelta / 100.
        mag_g = float(np.linalg.norm(np.ravel(g), ord=cg_kwargs.get("norm_ord", 1)))
        cg_resnorm = min(0.5, np.sqrt(mag_g)) * mag_g
        default_kwargs = {"absdelta": cg_absdelta, "resnorm": cg_resnorm, "norm_ord": 1, "_raise_nonposdef": False}
        cg_res = _cg.cg_re(lambda t, _p=pos: hessp(_p, t), g, **{**default_kwargs, **cg_kwargs})
        nat_g, info = cg_res.x, cg_res.info
        nhev += cg_res.nfev
        if info is not None and info < 0:
            raise ValueError("conjugate gradient failed")
        naive_ls_it = 0
        dd = nat_g
        grad_scaling = 1.
        for naive_ls_it in range(9):
            new_pos = pos - grad_scaling * dd
            new_energy, new_g = fun_and_grad(new_pos)
            nfev, njev = nfev + 1, njev + 1
            if new_energy <= energy:
                break
            grad_scaling /= 2
            if naive_ls_it == 5:
                gam = float(np.vdot(g, g))
                curv = float(np.vdot(g, hessp(pos, g)))
                nhev += 1
                grad_scaling = 1.
                dd = gam / curv * g
        else:
            grad_scaling = 0.
            status = -1
            break
        energy_diff = energy - new_energy
        old_fval = energy
        energy, pos, g = new_energy, new_pos, new_g
        descent_norm = grad_scaling * gradnorm(dd)
        if np.isnan(new_energy):
            raise ValueError("energy is NaN")
        min_cond = naive_ls_it < 2 and i > miniter
        if absdelta is not None and 0. <= energy_diff < absdelta and min_cond:
            status = 0
            break
        if descent_norm <= xtol and i > miniter:
            status = 0
            break
    else:
        status = i
    return OptimizeResults(pos, True, status, energy, g, i, nfev, njev, nhev)
