#!/usr/bin/env python3
"""Generate golden vectors from the REAL reference (NIFTy 8.5 numpy stack).

Run in the build container only (``/root/reference`` does not exist on the GPU
box)::

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

Writes ``tests/golden/*.npz``.  The reference is imported unmodified; the only
harness-side shim is ``np.asfarray`` (removed in NumPy 2, used by
``src/utilities.py:397``).  ``ducc0`` is absent, so the reference runs its own
``except ImportError`` branch (SciPy pocketfft, ``src/ducc_dispatch.py:88-92``).
Inputs come from ``np.random.default_rng(seed)`` so every consumer can rebuild them.
"""

import os
import sys

import numpy as np

np.asfarray = lambda a, dtype=np.float64: np.asarray(a, dtype=dtype)  # noqa: E731
sys.dont_write_bytecode = True
sys.path.insert(0, "/root/reference")
import nifty8 as ift  # noqa: E402
from nifty8 import ducc_dispatch as dd  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
CONVS = ("non_canonical_hartley", "canonical_hartley")


def gen_transforms():
    """``ducc_dispatch.hartley/fftn/ifftn`` (``src/ducc_dispatch.py:38-50``) incl. the
    shapes of ``test/test_ducc_scipy_fft.py:15-49`` and odd / partial-axes cases."""
    rng = np.random.default_rng(42)
    out = {}
    cases = [((10,), None), ((32, 32), None), ((16, 16, 16), None), ((12, 9), None),
             ((15, 27), None), ((8, 6, 10), None), ((16, 8, 4), (0, 1)), ((16, 8, 4), (2,)),
             ((6, 16, 8), (1,)), ((6, 16, 8), (1, 2)), ((8, 4, 16), (0, 2)), ((64,), None),
             ((1,), None), ((2, 2), None), ((7, 1, 5), None), ((128, 128), None)]
    for n, (shape, axes) in enumerate(cases):
        a = rng.normal(size=shape)
        out[f"h{n}_in"] = a
        out[f"h{n}_axes"] = np.array([-1] if axes is None else axes)
        for c in CONVS:
            ift.config.update("hartley_convention", c)
            out[f"h{n}_{c}"] = dd.hartley(a, axes=axes)
        z = rng.normal(size=shape) + 1j * rng.normal(size=shape)
        if z.size <= 2048:  # keep the fixture small; the c2c path is secondary
            out[f"h{n}_cin"] = z
            out[f"h{n}_fftn"] = dd.fftn(z, axes=axes)
            out[f"h{n}_ifftn"] = dd.ifftn(z, axes=axes)
    out["ncases"] = np.array(len(cases))
    ift.config.update("hartley_convention", CONVS[0])
    np.savez_compressed(os.path.join(OUT, "transforms.npz"), **out)


def gen_operator():
    """``HartleyOperator`` scaling / modes (``src/operators/harmonic_operators.py:200-217``)."""
    rng = np.random.default_rng(43)
    out = {}
    cases = [((16,), 0.3), ((8, 12), (0.5, 2.0)), ((4, 6, 8), 0.1)]
    for n, (shape, dist) in enumerate(cases):
        sp = ift.RGSpace(shape, distances=dist)
        op = ift.HartleyOperator(sp)
        x = rng.normal(size=shape)
        fx = ift.makeField(sp, x)
        y = op(fx)
        out[f"o{n}_in"] = x
        out[f"o{n}_dist"] = np.broadcast_to(np.asarray(dist, dtype=float), (len(shape),))
        out[f"o{n}_times"] = y.val
        out[f"o{n}_inverse"] = op.inverse_times(y).val
        out[f"o{n}_adjoint"] = op.adjoint_times(ift.makeField(op.target, x)).val
        z = x + 1j * rng.normal(size=shape)
        out[f"o{n}_cin"] = z
        out[f"o{n}_ctimes"] = op(ift.makeField(sp, z)).val
        fop = ift.FFTOperator(sp)
        out[f"o{n}_fft_times"] = fop(ift.makeField(sp, z)).val
        out[f"o{n}_fft_inverse"] = fop.inverse_times(ift.makeField(fop.target, z)).val
    out["ncases"] = np.array(len(cases))
    np.savez_compressed(os.path.join(OUT, "operator.npz"), **out)


CF_CASES = [
    # name, list of (shape, distances, fluct kwargs), offset_mean, offset_std
    ("cf1d", [((16,), 0.1, dict(fluctuations=(1., .1), loglogavgslope=(-1., .1),
                                flexibility=(1., .1), asperity=(.2, 2e-2)))], 0.0, (.1, .1)),
    ("cf2d", [((12, 12), 1. / 12, dict(fluctuations=(1e-1, 5e-3), loglogavgslope=(-1., 1e-2),
                                       flexibility=(1., .5), asperity=(.5, 5e-2)))], 0.0, (1e-3, 1e-4)),
    ("cf2d_odd", [((5, 6), (0.3, 0.5), dict(fluctuations=(3., 2.), loglogavgslope=(-2., 1.),
                                            flexibility=(3., 2.), asperity=None))], 1.5, (.1, .1)),
    ("cf3d", [((8, 8, 8), 1. / 8, dict(fluctuations=(1., .5), loglogavgslope=(-5., .2),
                                       flexibility=(1., .2), asperity=(.5, 5e-2)))], 2.0, (1e-1, 3e-2)),
    ("cfprod", [((8, 8), 1. / 8, dict(fluctuations=(1e-1, 5e-3), loglogavgslope=(-1., 1e-2),
                                      flexibility=(1., .5), asperity=(.5, 5e-2))),
                ((6,), 0.7, dict(fluctuations=(1., 1e-2), loglogavgslope=(-4., 1.),
                                 flexibility=(3.1, 1.), asperity=(.5, .1)))], 0.5, (1e-2, 1e-3)),
]


def build_cf(case):
    name, subs, off_mean, off_std = case
    cfm = ift.CorrelatedFieldMaker("")
    cfm.set_amplitude_total_offset(off_mean, off_std)
    for i, (shape, dist, kw) in enumerate(subs):
        pre = "" if len(subs) == 1 else f"ax{i}"
        cfm.add_fluctuations(ift.RGSpace(shape, dist), **kw, prefix=pre)
    return cfm, cfm.finalize(prior_info=0)


def gen_cf():
    """numpy-stack ``CorrelatedFieldMaker.finalize()`` operator
    (``src/library/correlated_fields.py:764-823``): value, Jacobian adjoint, tables."""
    out = {}
    for case in CF_CASES:
        name = case[0]
        for conv in CONVS:
            ift.config.update("hartley_convention", conv)
            cfm, cf = build_cf(case)
            rng = np.random.default_rng(7)
            pos = {k: rng.normal(size=cf.domain[k].shape) for k in sorted(cf.domain.keys())}
            npos = ift.MultiField.from_dict({k: ift.makeField(cf.domain[k], v) for k, v in pos.items()},
                                            cf.domain)
            val = cf(npos).val
            lin = cf(ift.Linearization.make_var(npos))
            ybar = rng.normal(size=val.shape)
            adj = lin.jac.adjoint_times(ift.makeField(cf.target, ybar))
            tang = {k: rng.normal(size=cf.domain[k].shape) for k in sorted(cf.domain.keys())}
            ntang = ift.MultiField.from_dict({k: ift.makeField(cf.domain[k], v) for k, v in tang.items()},
                                             cf.domain)
            jvp = lin.jac(ntang).val
            # full-latent metric of a Gaussian likelihood with diagonal inverse noise covariance w:
            # M t = J^T w J t + t  (SandwichOperator.make(J) with N^-1, + the prior's identity)
            wgt = np.exp(0.3 * rng.normal(size=val.shape))
            mt = lin.jac.adjoint_times(ift.makeField(cf.target, wgt * jvp)) + ntang
            tag = f"{name}_{conv}"
            for k, v in pos.items():
                out[f"{tag}_pos_{k}"] = v
                out[f"{tag}_tan_{k}"] = tang[k]
                out[f"{tag}_adj_{k}"] = adj[k].val
            out[f"{tag}_val"] = val
            out[f"{tag}_ybar"] = ybar
            out[f"{tag}_jvp"] = jvp
            out[f"{tag}_mweight"] = wgt
            for k in pos:
                out[f"{tag}_mt_{k}"] = mt[k].val
            out[f"{tag}_azm"] = np.asarray(cfm.azm.force(npos).val)
            for i, a in enumerate(cfm.get_normalized_amplitudes()):
                out[f"{tag}_namp{i}"] = a.force(npos).val
                out[f"{tag}_pindex{i}"] = a.target[0].pindex
        ift.config.update("hartley_convention", CONVS[0])
    np.savez_compressed(os.path.join(OUT, "correlated_field.npz"), **out)


def gen_cg():
    """``ift.ConjugateGradient`` (``src/minimization/conjugate_gradient.py:48-126``) on the
    xi-block metric ``R^T w R + 1`` with ``R = HT . diag(A)``; both controllers."""
    out = {}
    rng = np.random.default_rng(11)
    for n, (shape, dist) in enumerate([((16, 16), 1. / 16), ((8, 8, 8), 1. / 8), ((64,), 0.02)]):
        sp = ift.RGSpace(shape, dist)
        hsp = sp.get_default_codomain()
        ht = ift.HarmonicTransformOperator(hsp, sp)
        kl = hsp.get_k_length_array().val
        A = 1.0 / (1.0 + (kl / kl.max() * 6.0) ** 2)
        A[(0,) * len(shape)] = 1.0
        A = A * sp.total_volume ** 0.5 * 3.0
        w = np.exp(rng.normal(size=shape))
        R = ht @ ift.makeOp(ift.makeField(hsp, A))
        M = R.adjoint @ ift.makeOp(ift.makeField(sp, w)) @ R + ift.ScalingOperator(hsp, 1.)
        x0 = rng.normal(size=shape)
        b = R.adjoint(ift.makeField(sp, np.sqrt(w) * rng.normal(size=shape))).val + x0
        for cname, ctrl in (("absdelta", ift.AbsDeltaEnergyController(1e-6 * x0.size, iteration_limit=100)),
                            ("gradnorm", ift.GradientNormController(tol_rel_gradnorm=1e-4, iteration_limit=100)),
                            ("limit7", ift.AbsDeltaEnergyController(0., iteration_limit=7)),
                            ("reset", ift.GradientNormController(tol_abs_gradnorm=1e-9, iteration_limit=45))):
            en = ift.QuadraticEnergy(ift.makeField(hsp, x0), M, ift.makeField(hsp, b))
            res, status = ift.ConjugateGradient(ctrl)(en)
            tag = f"cg{n}_{cname}"
            out[f"{tag}_x"] = res.position.val
            out[f"{tag}_status"] = np.array(status)
            out[f"{tag}_nit"] = np.array(ctrl._itcount)
            out[f"{tag}_energy"] = np.array(res.value)
        out[f"cg{n}_A"], out[f"cg{n}_w"], out[f"cg{n}_x0"], out[f"cg{n}_b"] = A, w, x0, b
        out[f"cg{n}_dist"] = np.array(dist)
        out[f"cg{n}_Mx0"] = M(ift.makeField(hsp, x0)).val
    out["ncases"] = np.array(3)
    np.savez_compressed(os.path.join(OUT, "cg.npz"), **out)


if __name__ == "__main__":
    gen_transforms()
    gen_operator()
    gen_cf()
    gen_cg()
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)))
