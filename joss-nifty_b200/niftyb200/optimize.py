"""Newton-CG, the outer loop that calls the conjugate gradient.

Mirror of ``_newton_cg`` (``/root/reference/src/re/optimize.py:102-240``) on flat device tensors: same defaults,
same CG tolerances per Newton step (``:147-165``), same naive line search with the steepest-descent reset
(``:172-198``), same stopping tests and status codes (``:217-231``).  Host control flow, device-resident state:
per Newton step the host reads the energy (one scalar per function evaluation) and two norms.
"""

import datetime
from collections import namedtuple

import numpy as np
import torch

from .cg import _cg as _default_cg
from .transforms import vdot_device

OptimizeResults = namedtuple("OptimizeResults", "x success status fun jac nit nfev njev nhev",
                             defaults=(None, None, None, None, None, None))


def _norm(v, ord):
    return float(torch.linalg.vector_norm(v.reshape(-1).double(), ord=ord).cpu())


def newton_cg(fun_and_grad, x0, hessp, *, miniter=None, maxiter=None, energy_reduction_factor=0.1, old_fval=None,
              absdelta=None, norm_ord=None, xtol=1e-5, name=None, cg=None, cg_kwargs=None, time_threshold=None,
              custom_gradnorm=None):
    """``fun_and_grad(x) -> (energy: float, gradient tensor)``, ``hessp(x, t) -> tensor`` (or an object with
    ``at(x)`` returning an operator for the CG, which may then offer ``apply_with_curvature``)."""
    cg = _default_cg if cg is None else cg
    norm_ord = 1 if norm_ord is None else norm_ord                      # :121
    miniter = 0 if miniter is None else miniter
    maxiter = 200 if maxiter is None else maxiter
    xtol = xtol * x0.numel()                                            # :124
    pos = x0
    cg_kwargs = {} if cg_kwargs is None else cg_kwargs
    cg_name = name + "CG" if name is not None else None
    gradnorm = (lambda v: _norm(v, norm_ord)) if custom_gradnorm is None else custom_gradnorm
    energy, g = fun_and_grad(pos)                                       # :137
    energy = float(energy)
    nfev, njev, nhev = 1, 1, 0
    if np.isnan(energy):
        raise ValueError("energy is Nan")
    status, i = -1, 0
    for i in range(1, maxiter + 1):
        if old_fval and energy_reduction_factor:                        # :149-152
            cg_absdelta = energy_reduction_factor * (old_fval - energy)
        else:
            cg_absdelta = None if absdelta is None else absdelta / 100.
        mag_g = _norm(g, cg_kwargs.get("norm_ord", 1))                  # :153
        cg_resnorm = min(0.5, np.sqrt(mag_g)) * mag_g                   # :154-156
        default_kwargs = {"absdelta": cg_absdelta, "resnorm": cg_resnorm, "norm_ord": 1, "_raise_nonposdef": False,
                          "name": cg_name, "time_threshold": time_threshold}
        op = hessp.at(pos) if hasattr(hessp, "at") else (lambda t, out=None, _p=pos: hessp(_p, t))
        cg_res = cg(op, g, **{**default_kwargs, **cg_kwargs})           # :166
        nat_g, info = cg_res.x, cg_res.info
        nhev += cg_res.nfev
        if info is not None and info < 0:
            raise ValueError("conjugate gradient failed")
        naive_ls_it = 0
        dd = nat_g                                                      # negative descent direction
        grad_scaling = 1.
        ls_reset = False
        for naive_ls_it in range(9):                                    # :177-198
            new_pos = pos - grad_scaling * dd
            new_energy, new_g = fun_and_grad(new_pos)
            new_energy = float(new_energy)
            nfev, njev = nfev + 1, njev + 1
            if new_energy <= energy:
                break
            grad_scaling /= 2
            if naive_ls_it == 5:
                ls_reset = True
                gam = float(vdot_device(g, g).cpu())
                curv = float(vdot_device(g, op(g)).cpu())
                nhev += 1
                grad_scaling = 1.
                dd = gam / curv * g
        else:
            grad_scaling = 0.
            status = -1
            break
        energy_diff = energy - new_energy                               # :200-204
        old_fval = energy
        energy, pos, g = new_energy, new_pos, new_g
        descent_norm = grad_scaling * gradnorm(dd)
        if np.isnan(new_energy):
            raise ValueError("energy is NaN")
        min_cond = naive_ls_it < 2 and i > miniter                      # :219
        if absdelta is not None and 0. <= energy_diff < absdelta and min_cond:
            status = 0
            break
        if descent_norm <= xtol and i > miniter:
            status = 0
            break
        if time_threshold is not None and datetime.datetime.now() > time_threshold:
            status = i
            break
    else:
        status = i
    return OptimizeResults(x=pos, success=True, status=status, fun=energy, jac=g, nit=i, nfev=nfev, njev=njev, nhev=nhev)
