"""Parity cases shared by the CPU (emulator) and GPU test files.

Every function compares the niftyb200 path -- whichever native library the calling test's
fixture selected -- against the CPU oracle (``oracle/``, pinned to the reference through
``tests/golden``) or against the golden vectors directly.  Tolerances are the north-star
ones: rel-L2 <= 1e-12 (fp64) / 1e-5 (fp32) for applies, identical CG iteration counts.
"""

import sys

import numpy as np
import pytest
import torch

import niftyb200 as nb
from niftyb200.transforms import PointOp, device as nb_device, get_plan
from oracle import cg as ocg
from oracle import correlated_field as ocf
from oracle import metric as om
from oracle import transforms as otr

cgm = sys.modules["niftyb200.cg"]
CONVS = ("non_canonical_hartley", "canonical_hartley")
TOL = {np.float64: 1e-12, np.float32: 1e-5}


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64 if not np.iscomplexobj(a) else np.complex128), np.asarray(b)
    nrm = np.linalg.norm(b.ravel())
    return np.linalg.norm((a - b).ravel()) / (nrm if nrm > 0 else 1.0)


def dev(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a if dtype is None else a.astype(dtype)))
    return t.to(nb_device())


def host(t):
    return t.detach().cpu().numpy()


def clear_plans():
    """Developer knobs (NB2_TMA...) are read when a plan is created: drop the cached plans after changing one."""
    from niftyb200 import transforms
    transforms._plan_cache.clear()


# ---------------------------------------------------------------- transforms
def check_hartley_golden(golden):
    """ducc_dispatch.hartley/fftn/ifftn drop-ins == golden vectors made by the reference."""
    g = golden["transforms"]
    for n in range(int(g["ncases"])):
        a = g[f"h{n}_in"]
        axes = tuple(int(x) for x in g[f"h{n}_axes"])
        axes = None if axes == (-1,) else axes
        for c in CONVS:
            nb.config.update("hartley_convention", c)
            try:
                assert rel(nb.hartley(a, axes), g[f"h{n}_{c}"]) <= 1e-12, (n, c)
                out32 = nb.hartley(a.astype(np.float32), axes)
                assert out32.dtype == np.float32
                assert rel(out32, g[f"h{n}_{c}"]) <= 1e-5, (n, c)
            finally:
                nb.config.update("hartley_convention", CONVS[0])
        if f"h{n}_cin" in g:
            z = g[f"h{n}_cin"]
            assert rel(nb.fftn(z, axes), g[f"h{n}_fftn"]) <= 1e-12
            assert rel(nb.ifftn(z, axes), g[f"h{n}_ifftn"]) <= 1e-12


def check_hartley_shapes(shapes, dtypes=(np.float64, np.float32), seed=42):
    rng = np.random.default_rng(seed)
    for shape, axes in shapes:
        a64 = rng.normal(size=shape)
        for dt in dtypes:
            a = a64.astype(dt)
            for c in CONVS:
                ref = otr.hartley(a.astype(np.float64), axes, c)
                nb.config.update("hartley_convention", c)
                try:
                    got = nb.hartley(a, axes)
                finally:
                    nb.config.update("hartley_convention", CONVS[0])
                assert got.dtype == dt and got.shape == a.shape
                assert rel(got, ref) <= TOL[dt], (shape, axes, dt, c, rel(got, ref))


def check_product_hartley(shapes, dtypes=(np.float64, np.float32), seed=23):
    """``Plan.hartley(product_last=True)`` -- the joint three-pass transform of a (leading axes) x (last axis)
    product of spaces -- against the oracle's one transform per sub-space (correlated_fields.py:790-793), both
    conventions, plain and with the fused prologue / epilogue operands of the hot tail."""
    from niftyb200.transforms import PointOp, Plan
    rng = np.random.default_rng(seed)
    for shape in shapes:
        d = len(shape)
        a64 = rng.normal(size=shape)
        f64 = np.exp(0.3 * rng.normal(size=shape))
        m64, add64 = rng.normal(size=shape), rng.normal(size=shape)
        for dt in dtypes:
            tdt = torch.float64 if dt == np.float64 else torch.float32
            pl = Plan(shape, None, tdt)
            assert pl.is_fast
            a = a64.astype(dt)
            for c in CONVS:
                def ref_of(x):
                    lead = otr.hartley(x, tuple(range(d - 1)), c) if d > 1 else x
                    return otr.hartley(lead, (d - 1,), c)
                got = host(pl.hartley(dev(a), convention=c, product_last=True))
                ref = ref_of(a.astype(np.float64))
                assert rel(got, ref) <= TOL[dt], (shape, dt, c, rel(got, ref))
                # hot tail: field prologue, scale / offset / multiplier epilogue; adjoint: field + add + dot
                f, m, ad = f64.astype(dt), m64.astype(dt), add64.astype(dt)
                got = host(pl.hartley(dev(a), convention=c, product_last=True,
                                      pro=PointOp(2, dev(f), 0.7, amp_sym=False),
                                      epi=PointOp(scale=0.25, add_scalar=1.5, mul=dev(m))))
                ref = 0.25 * ref_of(0.7 * f.astype(np.float64) * a.astype(np.float64)) * m.astype(np.float64) + 1.5
                assert rel(got, ref) <= TOL[dt], (shape, dt, c, rel(got, ref))
                dot_out = torch.zeros(1, dtype=torch.float64, device=dev(a).device)
                got = host(pl.hartley(dev(a), convention=c, product_last=True,
                                      epi=PointOp(2, dev(f), 0.5, add=dev(ad), dot=dev(ad), dot_out=dot_out)))
                ref = 0.5 * f.astype(np.float64) * ref_of(a.astype(np.float64)) + ad.astype(np.float64)
                assert rel(got, ref) <= TOL[dt], (shape, dt, c, rel(got, ref))
                dref = float(np.vdot(ad.astype(np.float64), ref))
                assert abs(float(dot_out[0]) - dref) <= 10 * TOL[dt] * max(abs(dref), np.linalg.norm(ref)), (shape, dt)
                # the sub-space transforms commute, so the product transform is its own inverse up to N
                back = host(pl.hartley(pl.hartley(dev(a), convention=c, product_last=True), convention=c,
                                       product_last=True, epi=PointOp(scale=1.0 / a.size)))
                assert rel(back, a) <= 4 * TOL[dt]
            # |k|^2 tables belong to one isotropic space
            with pytest.raises(RuntimeError):
                pl.hartley(dev(a), product_last=True,
                           epi=PointOp(1, torch.ones(pl.ksq_size, dtype=tdt, device=dev(a).device), 1.0))
    gp = Plan((6, 10), None, torch.float64)
    assert not gp.is_fast
    with pytest.raises(RuntimeError):
        gp.hartley(dev(rng.normal(size=(6, 10))), product_last=True)


def make_product_pair(subs, tdt, kw=None):
    """(ours, oracle) correlated field on a product of sub-spaces ``[(shape, distances), ...]``."""
    kw = NP_KW if kw is None else kw
    mk, mo = nb.CorrelatedFieldMaker("", dtype=tdt), ocf.CorrelatedFieldMaker("")
    for m in (mk, mo):
        m.set_amplitude_total_offset(0.5, (1e-2, 1e-3))
        for i, (shape, dist) in enumerate(subs):
            m.add_fluctuations(shape, dist, **kw, prefix=f"ax{i}", non_parametric_kind="power")
    return mk.finalize(), mo.finalize()


def check_product_cf(subs, dt=np.float64, seed=9):
    """Correlated field on a product of two spaces whose joint grid is a power of two: the model runs on ONE
    joint plan (``cf.joint``); forward, xi-adjoint, full VJP and the metric against the oracle."""
    tdt = torch.float64 if dt == np.float64 else torch.float32
    cf, co = make_product_pair(subs, tdt)
    assert cf.joint is not None and cf.fuses_direction_update
    shape = cf.shape
    p = cf.init(seed)
    p["xi"] = p["xi"].astype(dt)
    tab = cf.tables(p)
    amps_o, azm_o = co.normalized_amplitudes(p), co.azm(p)
    rng = np.random.default_rng(seed + 1)
    for conv in CONVS:
        y = cf(p, convention=conv)
        assert rel(y, co(p, conv)) <= TOL[dt], (shape, dt, conv, rel(y, co(p, conv)))
        yb = rng.normal(size=shape).astype(dt)
        xb = host(cf.adjoint_xi(dev(yb), tab, convention=conv))
        ref = co.adjoint(amps_o, azm_o, yb.astype(np.float64), convention=conv)[0]
        assert rel(xb, ref) <= TOL[dt], (shape, dt, conv, rel(xb, ref))
    w = np.exp(0.3 * rng.normal(size=shape)).astype(dt)
    M = nb.CFMetric(cf, tab, dev(w))
    Mo = om.cf_metric_xi(co, amps_o, azm_o, w.astype(np.float64))
    x = rng.normal(size=shape).astype(dt)
    q, curv = M.apply_with_curvature(dev(x))
    qo = Mo(x.astype(np.float64))
    assert rel(host(q), qo) <= 4 * TOL[dt], (shape, dt, rel(host(q), qo))
    co_ = float(np.vdot(x.astype(np.float64), qo))
    assert abs(float(curv) - co_) <= 10 * TOL[dt] * abs(co_)
    # fused direction update d <- r + beta d inside the first pass
    r, d0 = rng.normal(size=shape).astype(dt), rng.normal(size=shape).astype(dt)
    d = dev(d0.copy())
    q2 = M.apply_with_curvature(d, pre=(dev(r), 0.37))[0]
    dref = r.astype(np.float64) + 0.37 * d0.astype(np.float64)
    assert rel(host(d), dref) <= TOL[dt]
    assert rel(host(q2), Mo(dref)) <= 4 * TOL[dt]


def check_vjp_reproducible(shape, dist, dt=np.float64, seed=31):
    """The spectrum cotangent (binned reduction in the row-pass epilogue, distributors.py:105-112) is accumulated
    in two-limb fixed point with integer atomics: bit-identical from run to run, equal (to rounding) to the fp64
    atomics of ``bin_out``, and a non-finite input poisons the bins instead of being dropped."""
    tdt = torch.float64 if dt == np.float64 else torch.float32
    cf, co = make_pair(shape, dist, tdt)
    p = cf.init(seed)
    tab = cf.tables(p)
    assert tab.amp_mode == 1 and cf.plans[0].is_fast
    rng = np.random.default_rng(seed)
    xi, yb = dev(rng.normal(size=shape).astype(dt)), dev(rng.normal(size=shape).astype(dt))
    runs = [cf.vjp(yb, tab, xi) for _ in range(3)]
    for r in runs[1:]:
        assert torch.equal(r[1][0], runs[0][1][0]) and r[2] == runs[0][2]
    bins = torch.zeros(cf.plans[0].ksq_size, dtype=torch.float64, device=xi.device)
    cf.plans[0].hartley(yb, epi=PointOp(tab.amp_mode, tab.amp, tab.azm * cf.dvol, bin_x=xi, bin_out=bins))
    ref = host(bins[cf.grids[0].unique_ksq.to(bins.device)])
    assert rel(host(runs[0][1][0]), ref) <= (1e-13 if dt == np.float64 else 1e-6)
    # tiny and huge magnitudes: the scale follows the data
    for fac in (1e-30, 1e30) if dt == np.float64 else (1e-12, 1e12):
        a = cf.vjp(yb * fac, tab, xi)[1][0]
        assert rel(host(a) / fac, ref) <= (1e-13 if dt == np.float64 else 1e-6), fac
    bad = yb.clone()
    bad.view(-1)[3] = float("nan")
    assert bool(torch.isnan(cf.vjp(bad, tab, xi)[1][0]).all())
    with pytest.raises(RuntimeError):
        cf.plans[0].hartley(yb, epi=PointOp(bin_x=xi, bin_fixed=torch.zeros(2 * cf.plans[0].ksq_size + 2, dtype=torch.int64,
                                                                             device=xi.device)))


def check_c2c_shapes(shapes, dtypes=(np.complex128, np.complex64), seed=17, expect_fast=None):
    """``fftn`` / ``ifftn`` drop-ins (ducc_dispatch.py:66-72) against the oracle (scipy pocketfft): complex and
    real input, forward unnormalised, backward scaled 1/N over the transformed axes."""
    rng = np.random.default_rng(seed)
    for shape, axes in shapes:
        z128 = rng.normal(size=shape) + 1j * rng.normal(size=shape)
        for dt in dtypes:
            z = z128.astype(dt)
            tol = 1e-12 if dt == np.complex128 else 1e-5
            if expect_fast is not None:
                launches0 = nb._lib.get().nb2_launch_count()
            f, b = nb.fftn(z, axes), nb.ifftn(z, axes)
            if expect_fast is not None:
                # fast path: <= 4 launches per transform (conj sweep + 2 strided + rows); the O(n^2) path
                # runs one NaiveDft launch per axis plus two copies -- and is told apart by the plan flag
                from niftyb200.transforms import _norm_axes
                rdt = torch.float32 if dt == np.complex64 else torch.float64
                assert bool(get_plan(z.shape, _norm_axes(z.ndim, axes), rdt).is_c2c_fast) == expect_fast, (shape, axes)
            assert f.dtype == dt and f.shape == z.shape
            assert rel(f, otr.fftn(z.astype(np.complex128), axes)) <= tol, (shape, axes, dt, "fftn")
            assert rel(b, otr.ifftn(z.astype(np.complex128), axes)) <= tol, (shape, axes, dt, "ifftn")
        # real input: complex result (harmonic_operators.py:110-117 passes real fields too)
        x = rng.normal(size=shape)
        assert rel(nb.fftn(x, axes), otr.fftn(x, axes)) <= 1e-12
        assert rel(nb.ifftn(x.astype(np.float32), axes), otr.ifftn(x.astype(np.float32).astype(np.float64), axes)) <= 1e-5


def check_inplace_and_roundtrip(shape, dt=np.float64):
    """in-place execution == out-of-place; H(H(x)) = N x (test_fft_operator.py:38-60)."""
    rng = np.random.default_rng(7)
    a = rng.normal(size=shape).astype(dt)
    x = dev(a)
    plan = get_plan(x.shape, None, x.dtype)
    y = plan.hartley(x)
    x2 = x.clone()
    plan.hartley(x2, out=x2)
    assert torch.equal(y, x2)
    assert torch.equal(x, dev(a)), "input modified by out-of-place transform"
    back = plan.hartley(y, epi=PointOp(scale=1.0 / a.size))
    assert rel(host(back), a) <= 10 * TOL[dt]


def check_operator_golden(golden):
    """HartleyOperator modes and volume factors (harmonic_operators.py:200-217)."""
    g = golden["operator"]
    for n in range(int(g["ncases"])):
        x, dist = g[f"o{n}_in"], g[f"o{n}_dist"]
        op = nb.HartleyOperator(nb.RGSpace(x.shape, tuple(dist)))
        y = op.times(x)
        assert rel(y, g[f"o{n}_times"]) <= 1e-12
        assert rel(op.inverse_times(y), g[f"o{n}_inverse"]) <= 1e-12
        assert rel(op.adjoint_times(x), g[f"o{n}_adjoint"]) <= 1e-12
        assert rel(op.times(g[f"o{n}_cin"]), g[f"o{n}_ctimes"]) <= 1e-12
        # FFTOperator (harmonic_operators.py:106-123): times on the position space, inverse_times from the target
        fop = nb.FFTOperator(nb.RGSpace(x.shape, tuple(dist)))
        z = g[f"o{n}_cin"]
        assert rel(fop.times(z), g[f"o{n}_fft_times"]) <= 1e-12
        assert rel(fop.inverse_times(z), g[f"o{n}_fft_inverse"]) <= 1e-12
        assert rel(fop.inverse_times(fop.times(z)), z) <= 1e-12                 # test_fft_operator.py:38-60
        assert np.iscomplexobj(fop.times(x)) and rel(fop.times(x), fop.times(x.astype(np.complex128))) <= 1e-13
        # adjointness <A a, b> == <a, A^T b> (extra.py:156-167), also of the operator on the harmonic partner
        for o2 in (fop, nb.FFTOperator(fop.target)):
            rng = np.random.default_rng(n)
            a, b = (rng.normal(size=x.shape) + 1j * rng.normal(size=x.shape) for _ in range(2))
            lhs, rhs = np.vdot(b, o2.times(a)), np.vdot(o2.adjoint_times(b), a)
            assert abs(lhs - rhs) <= 1e-12 * abs(lhs)
            assert rel(o2.adjoint_inverse_times(o2.adjoint_times(b)), b) <= 1e-12


def check_vdot():
    rng = np.random.default_rng(5)
    a, b = rng.normal(size=(37, 11)), rng.normal(size=(37, 11))
    assert abs(nb.vdot(a, b) - np.vdot(a, b)) <= 1e-12 * abs(np.vdot(a, b)) + 1e-13
    z, w = a + 1j * b, b - 2j * a
    assert abs(nb.vdot(z, w) - np.vdot(z, w)) <= 1e-12 * abs(np.vdot(z, w))
    i = np.arange(12)
    assert nb.vdot(i, i) == float(np.vdot(i, i))       # ints promoted to f64 (ducc_dispatch.py:82-85)
    a32, b32 = a.astype(np.float32), b.astype(np.float32)
    ref = np.vdot(a32.astype(np.float64), b32.astype(np.float64))
    assert abs(nb.vdot(a32, b32) - ref) <= 1e-12 * abs(ref) + 1e-12   # fp64 accumulation


# ---------------------------------------------------------------- correlated field
NP_KW = dict(fluctuations=(1., .5), loglogavgslope=(-3., .2), flexibility=(1., .2), asperity=(.5, 5e-2))


def make_pair(shape, dist, dtype=torch.float64, offset=2.0, kw=NP_KW, kind="power", offset_std=(1e-1, 3e-2)):
    mk = nb.CorrelatedFieldMaker("", dtype=dtype)
    mk.set_amplitude_total_offset(offset, offset_std)
    mk.add_fluctuations(shape, dist, **kw, non_parametric_kind=kind)
    mo = ocf.CorrelatedFieldMaker("")
    mo.set_amplitude_total_offset(offset, offset_std)
    mo.add_fluctuations(shape, dist, **kw, non_parametric_kind=kind)
    return mk.finalize(), mo.finalize()


def check_cf_golden(golden, names):
    """Full model == golden output of the reference numpy stack, both conventions."""
    from test_oracle_golden import CF_CASES, golden_pos, build_oracle_cf
    g = golden["correlated_field"]
    for name in names:
        subs, off_mean, off_std = CF_CASES[name]
        mk = nb.CorrelatedFieldMaker("")
        mk.set_amplitude_total_offset(off_mean, off_std)
        for i, (shape, dist, kw) in enumerate(subs):
            mk.add_fluctuations(shape, dist, **kw, prefix="" if len(subs) == 1 else f"ax{i}",
                                non_parametric_kind="power")
        cf = mk.finalize()
        co = build_oracle_cf(name)
        for conv in CONVS:
            tag = f"{name}_{conv}"
            p = golden_pos(g, tag, co)
            assert set(p) == set(cf.domain)
            y = cf(p, convention=conv)
            assert rel(y, g[f"{tag}_val"]) <= 1e-10, (tag, rel(y, g[f"{tag}_val"]))
            tab = cf.tables(p)
            xb = cf.adjoint_xi(dev(g[f"{tag}_ybar"]), tab, convention=conv)
            assert rel(host(xb), g[f"{tag}_adj_xi"]) <= 1e-10, tag


def check_cf_latent_vjp_golden(golden, names):
    """Jacobian adjoint to EVERY latent leaf (xi, zeromode, fluctuations, loglogavgslope, flexibility, asperity,
    spectrum; per sub-space for product spaces) == the reference's ``lin.jac.adjoint_times`` golden vectors
    (make_golden.py: ``adj_<leaf>``): hot-tail VJP + the amplitude chain's VJP kernel + the priors."""
    from test_oracle_golden import CF_CASES, golden_pos, build_oracle_cf
    g = golden["correlated_field"]
    for name in names:
        subs, off_mean, off_std = CF_CASES[name]
        mk = nb.CorrelatedFieldMaker("")
        mk.set_amplitude_total_offset(off_mean, off_std)
        for i, (shape, dist, kw) in enumerate(subs):
            mk.add_fluctuations(shape, dist, **kw, prefix="" if len(subs) == 1 else f"ax{i}", non_parametric_kind="power")
        cf, co = mk.finalize(), build_oracle_cf(name)
        assert all(f.on_kernel for f in cf.fl)
        for conv in CONVS:
            tag = f"{name}_{conv}"
            p = golden_pos(g, tag, co)
            amps, azm = cf.normalized_amplitudes(p)                 # the chain kernel
            for a, b in zip(amps, co.normalized_amplitudes(p)):
                assert rel(host(a), b) <= 1e-13, tag
            adj = cf.vjp_latent(g[f"{tag}_ybar"], p, convention=conv)
            assert set(adj) == set(cf.domain) | {"offset_mean"}
            for k in cf.domain:
                ref = g[f"{tag}_adj_{k}"]
                got = adj[k]
                got = host(got) if isinstance(got, torch.Tensor) else np.asarray(got)
                got = got.T if k.endswith("spectrum") else got           # numpy stack stores (2, nb - 2)
                # the zero-mode cotangent is a difference of two cancelling terms: 1e-9 instead of 1e-10
                assert rel(got, ref) <= (1e-9 if k.endswith("zeromode") else 1e-10), (tag, k, rel(got, ref))


def check_latent_metric_golden(golden, names):
    """The metric over the FULL latent vector, ``M t = J^T w J t + t`` (evi.py:108-109, likelihood.py:568-576),
    against vectors computed by the real reference (make_golden.py: ``mt_<leaf>`` = ``jac.adjoint_times(w *
    jac(t)) + t``); the JVP against ``jvp``; symmetry; and a full MGVI sample draw (``draw_linear_residual``,
    evi.py:83-144) whose solution must satisfy ``M x = j`` for the reference-built ``M``."""
    from test_oracle_golden import CF_CASES, golden_pos, build_oracle_cf
    g = golden["correlated_field"]
    for name in names:
        subs, off_mean, off_std = CF_CASES[name]
        mk = nb.CorrelatedFieldMaker("")
        mk.set_amplitude_total_offset(off_mean, off_std)
        for i, (shape, dist, kw) in enumerate(subs):
            mk.add_fluctuations(shape, dist, **kw, prefix="" if len(subs) == 1 else f"ax{i}", non_parametric_kind="power")
        cf, co = mk.finalize(), build_oracle_cf(name)
        for conv in CONVS:
            tag = f"{name}_{conv}"
            p, t = golden_pos(g, tag, co), golden_pos(g, tag, co, "tan")
            assert rel(host(cf.jvp_latent(t, p, convention=conv)), g[f"{tag}_jvp"]) <= 1e-12, tag
            M = nb.LatentMetric(cf, p, dev(g[f"{tag}_mweight"]), convention=conv)
            lay = M.layout
            tf = lay.flatten(t)
            mt = lay.unflatten(M(tf))
            for k in cf.domain:
                ref = g[f"{tag}_mt_{k}"]
                got = host(mt[k])
                got = got.T if k.endswith("spectrum") else got
                assert rel(got, ref) <= 1e-9, (tag, k, rel(got, ref))
            # symmetric: <u, M v> == <M u, v>
            rng = np.random.default_rng(1)
            u, v = dev(rng.normal(size=lay.size)), dev(rng.normal(size=lay.size))
            a, b = float(torch.dot(u, M(v))), float(torch.dot(M(u), v))
            assert abs(a - b) <= 1e-10 * max(abs(a), abs(b)), tag
        # MGVI sample on the full latent space: solve, then verify with the metric itself and its counts
        wd, wp = rng.normal(size=cf.shape), dev(rng.normal(size=lay.size))
        j = M.sample_rhs(dev(wd), wp)
        x, info = nb.draw_linear_residual(M, dev(wd), wp, cg_kwargs=dict(resnorm=1e-9 * float(torch.linalg.vector_norm(j)),
                                                                          maxiter=20 * lay.size))
        assert info == 0
        assert float(torch.linalg.vector_norm(M(x) - j) / torch.linalg.vector_norm(j)) <= 1e-8, name


def check_cf_latent_vjp_directional(shape, dist, kind="power", seed=5):
    """<ybar, d cf(p)[t]> == <vjp_latent(ybar), t> with the directional derivative by central differences of the
    ORACLE model (fp64), for a random direction t over all latent leaves -- at bin counts beyond the golden
    cases (the segmented scans of the chain kernels with many elements per thread)."""
    cf, co = make_pair(shape, dist, torch.float64, kind=kind)
    rng = np.random.default_rng(seed)
    p = cf.init(seed)
    t = {k: rng.standard_normal(np.shape(v)) for k, v in p.items()}
    ybar = rng.standard_normal(shape)
    adj = cf.vjp_latent(ybar, p)
    lhs = 0.0
    for k in cf.domain:
        a = adj[k]
        a = host(a) if isinstance(a, torch.Tensor) else np.asarray(a)
        lhs += float(np.vdot(a, t[k]))
    eps = 1e-6
    yp = co({k: p[k] + eps * t[k] for k in p})
    ym = co({k: p[k] - eps * t[k] for k in p})
    rhs = float(np.vdot(ybar, (yp - ym) / (2 * eps)))
    assert abs(lhs - rhs) <= 2e-7 * max(abs(lhs), abs(rhs)), (shape, kind, lhs, rhs)
    # and the kernel amplitudes against the oracle chain
    amps, _ = cf.normalized_amplitudes(p)
    for a, b in zip(amps, co.normalized_amplitudes(p)):
        assert rel(host(a), b) <= 1e-13


MATERN_KW = dict(scale=(1.0, 0.3), cutoff=(4.0, 1.0), loglogslope=(-4.0, 0.5))


def check_matern_latent_derivatives(shape, dist, kind="amplitude", renorm=True, seed=6):
    """Matern spectra (re/correlated_field.py:186-213) run the O(n_bins) chain with torch ops; its VJP / JVP come
    from autograd.  <ybar, d cf(p)[t]> == <vjp_latent(ybar), t> == <ybar, jvp_latent(t)> with the directional
    derivative by central differences of the ORACLE model, and the full-latent metric built on them is symmetric."""
    mk, mo = nb.CorrelatedFieldMaker("", dtype=torch.float64), ocf.CorrelatedFieldMaker("")
    for m in (mk, mo):
        m.set_amplitude_total_offset(0.5, (1e-1, 3e-2))
        m.add_fluctuations_matern(shape, dist, **MATERN_KW, renormalize_amplitude=renorm, non_parametric_kind=kind)
    cf, co = mk.finalize(), mo.finalize()
    rng = np.random.default_rng(seed)
    p = cf.init(seed)
    assert rel(cf(p), co(p)) <= 1e-12
    t = {k: rng.standard_normal(np.shape(v)) for k, v in p.items()}
    ybar = rng.standard_normal(shape)
    adj = cf.vjp_latent(ybar, p)
    lhs = 0.0
    for k in cf.domain:
        a = adj[k]
        a = host(a) if isinstance(a, torch.Tensor) else np.asarray(a)
        lhs += float(np.vdot(a, t[k]))
    eps = 1e-6
    yp = co({k: p[k] + eps * t[k] for k in p})
    ym = co({k: p[k] - eps * t[k] for k in p})
    rhs = float(np.vdot(ybar, (yp - ym) / (2 * eps)))
    assert abs(lhs - rhs) <= 2e-7 * max(abs(lhs), abs(rhs)), (shape, kind, lhs, rhs)
    jv = host(cf.jvp_latent(t, p))
    assert rel(jv, (yp - ym) / (2 * eps)) <= 1e-6
    mid = float(np.vdot(ybar, jv))
    assert abs(lhs - mid) <= 1e-10 * max(abs(lhs), abs(mid))
    # the full-latent metric on top of the chain derivatives
    lay = nb.LatentLayout(cf)
    M = nb.LatentMetric(cf, p, dev(np.exp(0.3 * rng.standard_normal(shape))))
    u, v = dev(rng.normal(size=lay.size)), dev(rng.normal(size=lay.size))
    a, b = float(torch.dot(u, M(v))), float(torch.dot(M(u), v))
    assert abs(a - b) <= 1e-10 * max(abs(a), abs(b))


def check_cf_vs_oracle(shape, dist, dt=np.float64, expect_mode=None, seed=3):
    tdt = torch.float64 if dt == np.float64 else torch.float32
    cf, co = make_pair(shape, dist, tdt)
    p = cf.init(seed)
    p["xi"] = p["xi"].astype(dt)
    tab = cf.tables(p)
    if expect_mode is not None:
        assert tab.amp_mode == expect_mode
    amps_o, azm_o = co.normalized_amplitudes(p), co.azm(p)
    for a, b in zip(tab.amps, amps_o):
        assert rel(host(a), b) <= 1e-12
    rng = np.random.default_rng(seed + 1)
    for conv in CONVS:
        y = cf(p, convention=conv)
        assert y.dtype == dt
        assert rel(y, co(p, conv)) <= TOL[dt], (shape, dt, conv, rel(y, co(p, conv)))
        yb = rng.normal(size=shape).astype(dt)
        xb = host(cf.adjoint_xi(dev(yb), tab, convention=conv))
        ref = co.adjoint(amps_o, azm_o, yb.astype(np.float64), convention=conv)[0]
        assert rel(xb, ref) <= TOL[dt], (shape, dt, conv, rel(xb, ref))
        # full cotangent incl. the spectrum (binned reduction, distributors.py:105-112)
        xi = p["xi"]
        for t in (tab, cf.streaming_tables(tab)):
            xb2, ampbars, azmbar, offbar = cf.vjp(dev(yb), t, dev(xi), convention=conv)
            rxb, rab, razm, roff = co.adjoint(amps_o, azm_o, yb.astype(np.float64), xi=xi.astype(np.float64),
                                              convention=conv)
            assert rel(host(xb2), rxb) <= TOL[dt]
            for a, b in zip(ampbars, rab):
                assert rel(host(a), b) <= 10 * TOL[dt], (shape, dt, conv, rel(host(a), b))
            scale = np.linalg.norm(np.concatenate([b.ravel() for b in rab])) + 1e-300
            assert abs(azmbar - razm) <= 10 * TOL[dt] * max(abs(razm), scale)
            assert abs(offbar - roff) <= 10 * TOL[dt] * max(abs(roff), np.linalg.norm(yb))
    # adjointness <A x, y> == <x, A^T y>  (extra.py:156-167)
    x = rng.normal(size=shape).astype(dt)
    Ax = host(cf.apply_xi(dev(x), tab, offset=0.0)).astype(np.float64)
    Aty = host(cf.adjoint_xi(dev(yb), tab)).astype(np.float64)
    lhs, rhs = np.vdot(Ax, yb.astype(np.float64)), np.vdot(x.astype(np.float64), Aty)
    assert abs(lhs - rhs) <= (1e-12 if dt == np.float64 else 1e-4) * max(abs(lhs), abs(rhs), 1e-30)


# 2-D Poissonian demo hyper-parameters (demos/0_intro.py:24-37)
C2_KW = dict(fluctuations=(1e-1, 5e-3), loglogavgslope=(-1., 1e-2), flexibility=(1., .5), asperity=(.5, 5e-2))


def metric_pair(shape, dist, dt=np.float64, seed=11, kw=NP_KW, likelihood="gaussian"):
    """(our metric, oracle metric, j, x0) of one synthetic sample draw (evi.py:83-144).

    ``likelihood="gaussian"``: dense per-voxel inverse variance ``w``; ``"poisson"``: the Poissonian
    metric with exp link, ``w = lambda = exp(cf(pos))`` (likelihood_impl.py:248-258), at the C2 demo
    hyper-parameters."""
    tdt = torch.float64 if dt == np.float64 else torch.float32
    if likelihood == "poisson":
        cf, co = make_pair(shape, dist, tdt, offset=0.0, kw=C2_KW, offset_std=(1e-3, 1e-4))
    else:
        cf, co = make_pair(shape, dist, tdt, kw=kw)
    p = cf.init(seed)
    amps_o, azm_o = co.normalized_amplitudes(p), co.azm(p)
    # both sides get the SAME amplitudes, so that they apply the same operator: the O(n_bins) amplitude
    # chain rounds differently in torch and NumPy (parallel vs sequential cumsum, 3e-13 at 512^2), which is
    # checked on its own in check_cf_vs_oracle and would otherwise be a (tiny) change of the linear system
    tab = cf.tables_from_amplitudes(amps_o, float(azm_o))
    rng = np.random.default_rng(seed + 1)
    if likelihood == "poisson":
        w = om.poisson_exp_weight(co(p)).astype(dt)
        rng.normal(size=shape)
    else:
        w = np.exp(0.3 * rng.normal(size=shape)).astype(dt)
    M = nb.CFMetric(cf, tab, dev(w))
    M._azm = tab.azm
    Mo = om.cf_metric_xi(co, amps_o, azm_o, w.astype(np.float64))
    Mo.parts = (co, amps_o, azm_o, w.astype(np.float64))
    x0 = rng.normal(size=shape).astype(dt)
    j = om.linear_sample_rhs(co, amps_o, azm_o, w.astype(np.float64), rng.normal(size=shape),
                             x0.astype(np.float64)).astype(dt)
    return M, Mo, j, x0


def M0_azm(M):
    return M._azm


def check_metric(shape, dist, dt=np.float64, likelihood="gaussian"):
    M, Mo, j, x0 = metric_pair(shape, dist, dt, likelihood=likelihood)
    t = dev(x0)
    q, curv = M.apply_with_curvature(t)
    ref = Mo(x0.astype(np.float64))
    assert rel(host(q), ref) <= TOL[dt]
    c_ref = np.vdot(x0.astype(np.float64), host(q).astype(np.float64))
    assert abs(float(curv.cpu()) - c_ref) <= 1e-12 * abs(c_ref)       # fused dot == dot of its own output
    assert rel(host(M(t)), ref) <= TOL[dt]
    M2 = nb.CFMetric(M.cf, M.cf.tables_from_amplitudes(M.tab.amps, float(M0_azm(M))), M.w, stream_amplitude=False)
    assert rel(host(M2(t)), ref) <= TOL[dt]          # gather (amp_mode 1) and streamed field agree
    # symmetric positive definite
    u, v = np.random.default_rng(0).normal(size=(2,) + tuple(shape)).astype(dt)
    a = np.vdot(host(M(dev(u))).astype(np.float64), v.astype(np.float64))
    b = np.vdot(u.astype(np.float64), host(M(dev(v))).astype(np.float64))
    assert abs(a - b) <= (1e-11 if dt == np.float64 else 1e-4) * max(abs(a), abs(b))


def check_sampled_metric(shape, dist, dt=np.float64, nsamples=3, likelihood="gaussian"):
    """SampledCFMetric == the mean over the samples of the oracle's per-sample metrics (optimize_kl.py:103-120),
    every sample with its own amplitude tables and weight; curvature; a CG solve on it with identical counts."""
    tdt = torch.float64 if dt == np.float64 else torch.float32
    if likelihood == "poisson":
        cf, co = make_pair(shape, dist, tdt, offset=0.0, kw=C2_KW, offset_std=(1e-3, 1e-4))
    else:
        cf, co = make_pair(shape, dist, tdt, kw=GENTLE_KW)
    rng = np.random.default_rng(21)
    tabs, ws, mats = [], [], []
    for s in range(nsamples):
        p = cf.init(100 + s)                          # expansion point of sample s: position + sample
        amps_o, azm_o = co.normalized_amplitudes(p), co.azm(p)
        tabs.append(cf.tables_from_amplitudes(amps_o, float(azm_o)))
        w = (om.poisson_exp_weight(co(p)) if likelihood == "poisson" else np.exp(0.3 * rng.normal(size=shape))).astype(dt)
        ws.append(dev(w))
        A = azm_o * co.outer_amplitude(amps_o)
        mats.append((A, w.astype(np.float64)))

    def Mo(t):
        acc = np.zeros(shape)
        for A, w in mats:
            h = co.harmonic_transform(A * t)
            g = w * h
            for gr, ax in reversed(list(zip(co.grids, co.axes))):
                g = (1.0 / gr.total_volume) * otr.hartley(g, axes=ax)
            acc += A * g
        return acc / len(mats) + t
    M = nb.SampledCFMetric(cf, tabs, ws)
    x = rng.normal(size=shape).astype(dt)
    ref = Mo(x.astype(np.float64))
    q, curv = M.apply_with_curvature(dev(x))
    assert rel(host(q), ref) <= TOL[dt], (shape, dt, rel(host(q), ref))
    c_ref = np.vdot(x.astype(np.float64), host(q).astype(np.float64))
    assert abs(float(curv.cpu()) - c_ref) <= 1e-12 * abs(c_ref)
    assert rel(host(M(dev(x))), ref) <= TOL[dt]
    # one sample == the plain metric
    M1 = nb.SampledCFMetric(cf, tabs[:1], ws[:1])
    assert rel(host(M1(dev(x))), host(nb.CFMetric(cf, tabs[0], ws[0])(dev(x)))) <= (1e-14 if dt == np.float64 else 1e-6)
    if dt == np.float64:
        j = rng.normal(size=shape)
        kw = dict(resnorm=1e-6 * float(np.linalg.norm(j)), maxiter=200)
        r, ro = cgm._cg(M, dev(j), None, **kw), ocg.cg_re(Mo, j, None, **kw)
        s_ = cgm._static_cg(M, dev(j), None, **kw)            # device-resident loop on the sampled metric
        assert abs(r.nit - ro.nit) <= 1 and r.info == ro.info == 0 and tuple(s_[1:]) == tuple(r[1:])
        assert rel(host(r.x), ro.x) <= 1e-5 and rel(Mo(host(r.x)), j) <= 1e-5


GENTLE_KW = dict(fluctuations=(1., .1), loglogavgslope=(-1.5, .1), flexibility=(1., .2), asperity=(.5, 5e-2))

def check_baseline_size(shape, dt):
    """Full oracle comparison at a BASELINE grid (SURVEY.md 8d "Parity gates"): plain Hartley in both
    conventions, CF forward, CF adjoint and the metric apply, rel-L2 <= 1e-12 (fp64) / 1e-5 (fp32; the
    oracle evaluated in fp64 on the fp32-rounded inputs).  The oracle runs on all host cores."""
    import os
    otr.set_nthreads(os.cpu_count() or 1)
    try:
        tdt = torch.float64 if dt == np.float64 else torch.float32
        dist = tuple(1.0 / s for s in shape)
        cf, co = make_pair(shape, dist, tdt)
        p = cf.init(3)
        p["xi"] = p["xi"].astype(dt)
        xi64 = p["xi"].astype(np.float64)
        amps_o, azm_o = co.normalized_amplitudes(p), co.azm(p)
        own = cf.tables(p)                     # our O(n_bins) amplitude chain (torch; parallel scan)
        out = {"amplitude chain": max(rel(host(a), b) for a, b in zip(own.amps, amps_o))}
        # the hot path's inputs are the amplitude tables: the applies are compared on the SAME tables (the
        # chain's own rounding, 3e-13 at 141k bins, would otherwise enter the metric twice); the full model
        # with our own chain is the "forward (full model)" entry
        tab = cf.tables_from_amplitudes(amps_o, float(azm_o))
        x = dev(p["xi"])
        plan = get_plan(shape, None, tdt)
        out["forward (full model)"] = rel(host(cf.apply_xi(x, own)), co(p))
        del own
        for conv in CONVS:
            got = host(plan.hartley(x, convention=conv))
            out["hartley " + conv] = rel(got, otr.hartley(xi64, None, conv))
            del got
        out["forward"] = rel(host(cf.apply_xi(x, tab)), co(p))
        rng = np.random.default_rng(4)
        yb = rng.standard_normal(shape).astype(dt)
        out["adjoint"] = rel(host(cf.adjoint_xi(dev(yb), tab)), co.adjoint(amps_o, azm_o, yb.astype(np.float64))[0])
        w = np.exp(0.3 * rng.standard_normal(shape)).astype(dt)
        M = nb.CFMetric(cf, tab, dev(w))
        Mo = om.cf_metric_xi(co, amps_o, azm_o, w.astype(np.float64))
        q, curv = M.apply_with_curvature(x)
        ref = Mo(xi64)
        out["metric"] = rel(host(q), ref)
        out["curvature"] = abs(float(curv.cpu()) - np.vdot(xi64, ref)) / abs(np.vdot(xi64, ref))
        bad = {k: v for k, v in out.items() if not v <= TOL[dt]}
        assert not bad, (shape, dt, out)
        return out
    finally:
        otr.set_nthreads(1)


# ---------------------------------------------------------------- CG iteration-count parity
# north_star: "an identical CG iteration count to the same residual tolerance".
#
# What makes this delicate (measured on the ORACLE ALONE): the metric of a correlated
# field has an outlying, tightly clustered spectrum, and the CG recurrence passes through near-breakdown
# steps (a step whose energy decrease is 10-100x smaller than its neighbours').  Around such a step a
# 1e-16 relative rounding difference is amplified to 1e-3 of the energy for two or three iterations and
# then decays back to 1e-16 (the iterates re-converge): the reference run against a copy of itself in a
# second arithmetic shows exactly this.  A stopping threshold that happens to be crossed inside such a
# burst is decided by rounding in the reference too.
#
# Therefore two kinds of checks:
#  * STRICT (`check_cg_strict`): thresholds read off the oracle's own convergence history such that the
#    stopping test is decided with a margin at every iteration it is evaluated (threshold a factor 1.3 away
#    from the tested values and NOISE_MARGIN times the reference's own rounding sensitivity away from them).
#    For these rules nit / info / nfev must be IDENTICAL.
#  * demo-style thresholds (`check_cg_parity`): identical counts, or -- recorded in the assertion message --
#    a stopping iteration that is CONSISTENT with the oracle's own history and its measured per-iteration
#    rounding sensitivity (`acceptable_stops`): the reference could not have stopped earlier and may stop
#    there when each tested value is moved by at most ENVELOPE times the deviation that the oracle shows
#    when it is run in a second arithmetic (`other_arithmetic`: another summation order; `noisy_metric`:
#    rounding-sized transform noise).  Outside the bursts that sensitivity is ~1e-12 of the value, so the
#    criterion forces equality there; it is only loose where the reference itself is undetermined.
NOISE_MARGIN = 50.0
ENVELOPE = 3.0
WINDOW_SEEDS = (1, 2, 3, 4, 5, 6)

# Size of the rounding noise injected into the oracle to measure its sensitivity (relative L2, white, per
# harmonic transform).  Measured accuracy of one transform against an extended-precision one at 512^2:
# pocketfft 3e-16, ours 8e-16 (fp64) / 3e-7 (fp32); two CG implementations also differ in every vector
# update and reduction (FMA contraction, summation order), which measured on the kernel emulator is
# equivalent to about ten times the transform noise -- hence 1e-14 / 1e-6.
TRANSFORM_EPS = {np.float64: 1e-14, np.float32: 1e-6}


class other_arithmetic:
    """Context: the oracle's inner products summed in another order (pairwise ``np.sum`` of the products
    instead of the BLAS dot) -- every reduction of a second implementation rounds differently, and the
    reference recurrence is as sensitive to that as to the transform noise."""

    def __enter__(self):
        self.saved = ocg._vdot
        ocg._vdot = lambda a, b: float(np.sum(np.ravel(a).astype(np.float64) * np.ravel(b).astype(np.float64)))

    def __exit__(self, *a):
        ocg._vdot = self.saved


def noisy_metric(Mo, dt, seed):
    """The oracle metric `A H w H A + 1` with rounding-sized noise (TRANSFORM_EPS, white) added to the output
    of either harmonic transform -- a model of "the same operator evaluated by another FFT"; for fp32 the
    result is also rounded to fp32."""
    co, amps, azm, w = Mo.parts
    A = azm * co.outer_amplitude(amps)
    rng = np.random.default_rng(seed)
    eps = TRANSFORM_EPS[dt]

    def H(x):
        y = co.harmonic_transform(x)
        return y + (eps * np.sqrt(np.mean(y * y))) * rng.normal(size=y.shape)

    def mat(t):
        t64 = t.astype(np.float64)
        return (A * H(w * H(A * t64)) + t64).astype(dt)
    return mat


def count_window(run_oracle, Mo):
    """Iteration counts of the oracle itself in a second arithmetic (`run_oracle(mat) -> result`)."""
    nits = [run_oracle(Mo).nit]
    with other_arithmetic():
        nits += [run_oracle(noisy_metric(Mo, np.float64, sd)).nit for sd in WINDOW_SEEDS]
    return nits


def oracle_histories(Mo, j, x0, nmax, dt=np.float64, nseeds=2):
    """Energy-decrease / residual-norm histories of `nmax` forced iterations of the fp64 oracle, plus the
    largest deviation from them of `nseeds` runs of the oracle on `noisy_metric` (with all vectors stored in
    `dt`): the reference recurrence's own sensitivity to rounding of that size, per iteration.  Both
    reference stacks run the same recurrence on the same energy (quadratic_energy.py:35-37 ==
    re/conjugate_gradient.py:191), so one set of histories serves both."""
    def run(mat, jj, xx):
        h, g = [], []
        r0 = mat(xx) - jj
        e0 = ocg._vdot((r0 - jj) / 2, xx)
        ocg.cg_re(mat, jj, xx, resnorm=0.0, maxiter=nmax, _raise_nonposdef=False, history=h, res_history=g)
        return np.array([e0] + h), np.array(g)
    h, g = run(Mo, j, x0)
    dE = h[:-1] - h[1:]                       # dE[i-1]: energy decrease of iteration i
    nE, nG = np.zeros_like(dE), np.zeros_like(g)
    for sd in range(1, nseeds + 1):
        with other_arithmetic():
            hp, gp = run(noisy_metric(Mo, dt, sd), j.astype(dt), x0.astype(dt))
        m = min(len(hp), len(h)) - 1
        nE[:m] = np.maximum(nE[:m], np.abs((hp[:m] - hp[1:m + 1]) - dE[:m]))
        nG[:m] = np.maximum(nG[:m], np.abs(gp[:m] - g[:m]))
        nE[m:], nG[m:] = np.inf, np.inf       # a noisy run that broke down early: nothing known beyond
    return dE, g, nE, nG


def separated_thresholds(vals, noise, miniter, count=3, gap=1.3):
    """Thresholds t for the rule "stop at the first iteration i >= miniter with vals[i] < t" such that every
    comparison the rule makes is decided with a margin: |vals[k] - t| >= NOISE_MARGIN * noise[k] for
    miniter <= k <= i (noise = the reference's own rounding sensitivity at that iteration) and t lies a
    factor >= `gap` away from vals[i] and from every earlier value (iterations are 1-based: vals[i-1]
    belongs to iteration i).  Returns up to `count` pairs (threshold, stopping iteration) spread over the
    history: the earliest, the latest and the ones in between."""
    cands = []
    for i in range(miniter + 1, len(vals) + 1):
        prev = vals[miniter - 1:i - 1]
        v = vals[i - 1]
        if not (v > 0 and v * gap * gap <= prev.min()):
            continue
        t = float(np.sqrt(v * prev.min()))
        k = np.arange(miniter - 1, i)
        if np.all(np.abs(vals[k] - t) >= NOISE_MARGIN * noise[k]):
            cands.append((t, i))
    if len(cands) <= count:
        return cands
    pick = sorted(set(int(round(x)) for x in np.linspace(0, len(cands) - 1, count)))
    return [cands[q] for q in pick]


def check_cg_strict(shape, dist, dt=np.float64, likelihood="gaussian", nmax=40, min_checks=2):
    """IDENTICAL nit / info / nfev of `_cg`, `_static_cg` and `ConjugateGradient` for stopping rules with
    separated thresholds (absdelta and resnorm / energy and gradient-norm controllers).  fp32: our solver
    runs in fp32, the oracle in fp64 on the same fp32-rounded inputs; the thresholds are
    then separated with respect to fp32-sized rounding noise.  Returns the number of solves compared."""
    M, Mo, j, x0 = metric_pair(shape, dist, dt, kw=GENTLE_KW, likelihood=likelihood)
    jt, xt = dev(j), dev(x0)
    j64, x64 = j.astype(np.float64), x0.astype(np.float64)
    miniter = 6 if dt == np.float64 else 3
    xtol, etol = (1e-4, 1e-5) if dt == np.float64 else (1e-2, 1e-3)
    checked = 0
    dE, g, nE, nG = oracle_histories(Mo, j64, x64, nmax, dt)
    rules = [dict(absdelta=t, miniter=miniter, maxiter=nmax) for t, _ in separated_thresholds(dE, nE, miniter)]
    rules += [dict(resnorm=t, miniter=miniter, maxiter=nmax) for t, _ in separated_thresholds(g, nG, miniter)]
    for kw in rules:
        ro = ocg.cg_re(Mo, j64, x64, **kw)
        r = cgm._cg(M, jt, xt, **kw)
        assert (r.nit, r.info, r.nfev) == (ro.nit, ro.info, ro.nfev), ("_cg", shape, dt, kw, tuple(r[1:]), tuple(ro[1:]))
        so = ocg.static_cg_re(Mo, j64, x64, **kw)
        st = cgm._static_cg(M, jt, xt, **kw)
        assert (st.nit, st.info, st.nfev) == (so.nit, so.info, so.nfev), ("_static_cg", shape, dt, kw, tuple(st[1:]), tuple(so[1:]))
        assert rel(host(r.x), ro.x) <= xtol and rel(host(st.x), so.x) <= xtol, (shape, dt, kw, rel(host(r.x), ro.x))
        checked += 2
    # numpy stack: AbsDeltaEnergyController looks at |E_old - E| from the first step on,
    # GradientNormController at |grad|; neither has a minimum iteration count
    ctrls = [("AbsDeltaEnergyController", dict(deltaE=t, iteration_limit=nmax)) for t, _ in separated_thresholds(dE, nE, 1)]
    ctrls += [("GradientNormController", dict(tol_abs_gradnorm=t, iteration_limit=nmax))
              for t, _ in separated_thresholds(g, nG, 1)]
    for name, kw in ctrls:
        eo, so, nito = ocg.conjugate_gradient_np(ocg.QuadraticEnergy(x64, Mo, j64), getattr(ocg, name)(**kw))
        solver = nb.ConjugateGradient(getattr(nb, name)(**kw))
        e, status = solver(nb.QuadraticEnergy(xt, M, jt))
        assert (solver.nit, status) == (nito, so), ("ConjugateGradient", shape, dt, name, kw, solver.nit, nito)
        assert abs(e.value - eo.value) <= etol * abs(eo.value), (shape, dt, name, kw, e.value, eo.value)
        checked += 1
    assert checked >= min_checks, ("too few separated stopping rules", shape, dt, likelihood, rules, ctrls)
    return checked


def acceptable_stops(vals, noise, t, miniter, maxiter):
    """Stopping iterations of the rule "first i >= miniter with vals[i] < t, else maxiter" that are consistent
    with the oracle history `vals` when every value may move by ENVELOPE * `noise` (iterations 1-based):
    from the first iteration whose lower envelope is below `t` up to the first whose upper envelope is."""
    acc = set()
    lo, hi = vals - ENVELOPE * noise, vals + ENVELOPE * noise
    for i in range(max(miniter, 1), min(maxiter, len(vals)) + 1):
        if lo[i - 1] < t:
            acc.add(i)
        if hi[i - 1] < t:
            return acc
    acc.add(maxiter)
    return acc


def _same_counts(r, ro, kw, strict, hist, j):
    """Identical (nit, info, nfev); otherwise (demo-style thresholds that fall inside a near-breakdown burst,
    see above) a stopping iteration consistent with the oracle's history and sensitivity (`acceptable_stops`),
    which the assertion message records.  `strict`: identical, no exception."""
    if (r.nit, r.info, r.nfev) == (ro.nit, ro.info, ro.nfev):
        return
    assert not strict, (kw, tuple(r[1:]), tuple(ro[1:]))
    dE, g, nE, nG = hist
    maxiter = kw.get("maxiter", 200)
    miniter = kw.get("miniter", min(6, maxiter))
    if kw.get("absdelta") is not None:
        acc = acceptable_stops(dE, nE, kw["absdelta"], miniter, maxiter)
    else:
        acc = acceptable_stops(g, nG, kw.get("resnorm", 1e-5 * float(np.linalg.norm(j))), miniter, maxiter)
    assert r.nit in acc and r.info in (0, r.nit), \
        f"count {r.nit} (reference {ro.nit}) is not a stopping iteration consistent with the reference: {sorted(acc)} for {kw}"


def check_cg_parity(shape, dist, strict=True):
    """Demo-style stopping rules (absdelta ~ 1e-4 * size / 10, 0_intro.py:203-206; the library defaults):
    identical iteration counts / info codes and matching solutions vs the reference CGs.

    ``strict``: identical counts, energy histories to 1e-9 and solutions to 1e-8 (grids of <= 4096 points,
    where no near-breakdown burst precedes the stopping iteration).  Otherwise see `_same_counts`; the
    histories are then compared to 1e-12 over the first five iterations and to 2e-2 inside the bursts."""
    htol, xtol = (1e-9, 1e-8) if strict else (2e-2, 2e-2)      # 2e-2: a stop inside a near-breakdown burst
    M, Mo, j, x0 = metric_pair(shape, dist, np.float64, kw=GENTLE_KW)
    jt, xt = dev(j), dev(x0)
    n = j.size

    hist = None if strict else oracle_histories(Mo, j, x0, 200)
    for kw in (dict(absdelta=1e-5 * n, maxiter=100), dict(resnorm=1e-4 * np.linalg.norm(j), maxiter=100),
               dict(absdelta=1e-12, maxiter=27), dict()):
        h1, h2 = [], []
        r = cgm._cg(M, jt, xt, history=h1, **kw)
        ro = ocg.cg_re(Mo, j, x0, history=h2, **kw)
        _same_counts(r, ro, kw, strict, hist, j)
        m = min(len(h1), len(h2))
        np.testing.assert_allclose(h1[:m], h2[:m], rtol=htol)
        np.testing.assert_allclose(h1[:min(m, 5)], h2[:min(m, 5)], rtol=1e-12)     # before the first burst
        if r.nit == ro.nit:
            assert rel(host(r.x), ro.x) <= xtol
        s = cgm._static_cg(M, jt, xt, **kw)
        so = ocg.static_cg_re(Mo, j, x0, **kw)
        _same_counts(s, so, kw, strict, hist, j)
        if s.nit == so.nit:
            assert rel(host(s.x), so.x) <= xtol
    # x0 = None branch
    r, ro = cgm._cg(M, jt, None, absdelta=1e-7 * n), ocg.cg_re(Mo, j, None, absdelta=1e-7 * n)
    if strict:
        assert (r.nit, r.info, r.nfev) == (ro.nit, ro.info, ro.nfev)
    if r.nit == ro.nit:
        assert rel(host(r.x), ro.x) <= xtol
    # numpy-stack interface
    g0 = float(np.linalg.norm(Mo(x0) - j))          # gradient norm at the start (tol_rel_gradnorm reference)
    for mk, rule in ((lambda m: m.AbsDeltaEnergyController(1e-6 * n, iteration_limit=100), ("dE", 1e-6 * n, 100)),
                     (lambda m: m.GradientNormController(tol_rel_gradnorm=1e-4, iteration_limit=100), ("g", 1e-4 * g0, 100)),
                     (lambda m: m.AbsDeltaEnergyController(0., iteration_limit=23), ("dE", 0.0, 23))):
        solver = nb.ConjugateGradient(mk(nb))
        e, status = solver(nb.QuadraticEnergy(xt, M, jt))
        eo, so, nito = ocg.conjugate_gradient_np(ocg.QuadraticEnergy(x0, Mo, j), mk(ocg))
        assert status == so
        if solver.nit != nito:
            assert not strict, (solver.nit, nito)
            dE, g, nE, nG = hist
            vals, noise = (dE, nE) if rule[0] == "dE" else (g, nG)
            acc = acceptable_stops(vals, noise, rule[1], 1, rule[2])
            assert solver.nit in acc, \
                f"count {solver.nit} (reference {nito}) is not consistent with the reference: {sorted(acc)} for {rule}"
        else:
            assert rel(host(e.position), eo.position) <= xtol
            assert abs(e.value - eo.value) <= (1e-10 if strict else htol) * abs(eo.value)


def check_static_cg_device_resident(shape, dist, dt=np.float64, likelihood="gaussian"):
    """The device-resident ``static_cg`` (scalars, stopping tests, alpha / beta on the device; iterations enqueued
    ahead of the decision and gated) against the host-driven loop of the same kernels: identical nit / info /
    nfev and the same solution, for every stopping rule, across the exact-residual step at iteration 20, for
    several polling intervals, with x0 = None, and for the failure codes."""
    M, Mo, j, x0 = metric_pair(shape, dist, dt, kw=GENTLE_KW, likelihood=likelihood)
    jt, xt = dev(j), dev(x0)
    n = j.size
    tol = 1e-12 if dt == np.float64 else 1e-5
    for kw in (dict(absdelta=1e-5 * n, maxiter=100), dict(resnorm=1e-4 * float(np.linalg.norm(j)), maxiter=100),
               dict(absdelta=1e-12, maxiter=27), dict(), dict(absdelta=0.0, miniter=3, maxiter=41)):
        h = cgm._static_cg(M, jt, xt, device_resident=False, **kw)
        for poll in (1, 4, 7):
            s = cgm._static_cg(M, jt, xt, poll=poll, **kw)
            assert tuple(s[1:]) == tuple(h[1:]), (shape, dt, kw, poll, tuple(s[1:]), tuple(h[1:]))
            assert rel(host(s.x), host(h.x)) <= tol, (shape, dt, kw, poll)
    s, h = cgm._static_cg(M, jt, None, absdelta=1e-7 * n), cgm._static_cg(M, jt, None, absdelta=1e-7 * n, device_resident=False)
    assert tuple(s[1:]) == tuple(h[1:]) and rel(host(s.x), host(h.x)) <= tol
    # the public entry point takes the device-resident path by itself for a CFMetric
    x, info = nb.static_cg(M, jt, xt, absdelta=1e-5 * n, maxiter=100)
    assert info == 0 and rel(host(x), host(cgm._static_cg(M, jt, xt, absdelta=1e-5 * n, maxiter=100, device_resident=False).x)) <= tol

    class Neg:          # not positive definite: info = -1 after the first step (:272-276), on the device
        device_resident_cg = True

        def apply_with_curvature(self, t, out=None, pre=None, curv_out=None):
            out.copy_(-t)
            curv_out.copy_(-(t.double() * t.double()).sum().reshape(1))
            return out, curv_out

        def __call__(self, t, out=None):
            return -t
    r = cgm._static_cg(Neg(), jt, None)
    assert (r.info, r.nit) == (-1, 1)


def check_newton_cg_parity(n=64, seed=2):
    """``newton_cg`` against the oracle's restatement of ``_newton_cg`` (optimize.py:102-240) on analytic convex
    functions: identical nit / status / nfev / njev / nhev and the same minimiser -- the control flow (CG tolerances
    per step, line search with the steepest-descent reset, stopping tests)."""
    from oracle import optimize as oopt
    rng = np.random.default_rng(seed)
    a, b, d = rng.uniform(0.5, 2.0, n), rng.normal(size=n), rng.uniform(0.1, 3.0, n)
    at, bt, dt_ = dev(a), dev(b), dev(d)

    def f_np(x):
        return float(np.sum(np.cosh(a * x)) - b @ x + 0.5 * np.sum(d * x * x)), a * np.sinh(a * x) - b + d * x

    def h_np(x, t):
        return (a * a * np.cosh(a * x) + d) * t

    def f_t(x):
        return float((torch.cosh(at * x).sum() - bt @ x + 0.5 * (dt_ * x * x).sum()).cpu()), at * torch.sinh(at * x) - bt + dt_ * x

    def h_t(x, t):
        return (at * at * torch.cosh(at * x) + dt_) * t
    x0 = rng.normal(size=n) * 2.0
    for kw in (dict(), dict(absdelta=1e-8, miniter=2), dict(xtol=1e-9, maxiter=7), dict(energy_reduction_factor=None, absdelta=1e-6)):
        r = nb.newton_cg(f_t, dev(x0), h_t, **kw)
        ro = oopt.newton_cg(f_np, x0, h_np, **kw)
        assert (r.nit, r.status, r.nfev, r.njev, r.nhev) == (ro.nit, ro.status, ro.nfev, ro.njev, ro.nhev), (kw, r[2:], ro[2:])
        # inexact inner solves: the iterates agree to the rounding of the two stacks amplified by the CG
        # (a run cut off mid-descent by maxiter is the least converged: 2e-7 in x)
        assert rel(host(r.x), ro.x) <= 1e-5 and abs(r.fun - ro.fun) <= 1e-6 * abs(ro.fun), (kw, rel(host(r.x), ro.x), r.fun - ro.fun)


def check_gaussian_model_and_geovi(shape, dist, seed=4):
    """GaussianCFModel: energy against the oracle model, gradient by central differences of the oracle energy,
    a MAP fit with newton_cg (energy decreases to a stationary point), the sample-averaged KL, and the geoVI
    update (evi.py:201-266): its solution must zero the residual it minimises, and ``maxiter=0`` must skip."""
    cf, co = make_pair(shape, dist, torch.float64, kw=GENTLE_KW)
    rng = np.random.default_rng(seed)
    truth = cf.init(seed)
    sig = co(truth)
    w = np.full(shape, 1.0 / 0.1 ** 2)
    data = sig + 0.1 * rng.normal(size=shape)
    model = nb.GaussianCFModel(cf, dev(data), dev(w))
    lay = model.layout

    def e_oracle(p):
        r = data - co(p)
        return 0.5 * float(np.sum(w * r * r)) + 0.5 * sum(float(np.sum(np.asarray(v) ** 2)) for v in p.values())
    p0 = cf.init(seed + 1)
    x = lay.flatten(p0)
    e, g = model.energy_and_grad(x)
    assert abs(e - e_oracle(p0)) <= 1e-11 * abs(e)
    t = {k: rng.standard_normal(np.shape(v)) for k, v in p0.items()}
    eps = 1e-6
    fd = (e_oracle({k: p0[k] + eps * t[k] for k in p0}) - e_oracle({k: p0[k] - eps * t[k] for k in p0})) / (2 * eps)
    an = float(torch.dot(g, lay.flatten(t)))
    assert abs(fd - an) <= 1e-6 * max(abs(fd), abs(an)), (fd, an)
    # MAP
    res = nb.newton_cg(model.energy_and_grad, x, model.metric_at and (lambda pos, tt: model.metric_at(pos)(tt)), maxiter=15,
                       absdelta=1e-3)
    assert res.fun < e and res.status >= 0
    e2, g2 = model.energy_and_grad(res.x)
    assert abs(e2 - res.fun) <= 1e-10 * abs(e2)
    # MGVI sample at the MAP point, its mirror, the KL value / gradient / metric over the pair
    M = model.metric_at(res.x)
    wd, wp = dev(rng.normal(size=shape)), dev(rng.normal(size=lay.size))
    smpl, info = nb.draw_linear_residual(M, wd, wp, cg_kwargs=dict(resnorm=1e-8 * float(torch.linalg.vector_norm(M.sample_rhs(wd, wp))),
                                                                 maxiter=10 * lay.size))
    assert info == 0
    samples = [smpl, -smpl]
    ek, gk = model.kl_energy_and_grad(samples)(res.x)
    ea, ga = model.energy_and_grad(res.x + smpl)
    eb, gb = model.energy_and_grad(res.x - smpl)
    assert abs(ek - 0.5 * (ea + eb)) <= 1e-12 * abs(ek) and rel(host(gk), host(0.5 * (ga + gb))) <= 1e-12
    tt = dev(rng.normal(size=lay.size))
    mk = model.kl_metric(samples).at(res.x)(tt)
    assert rel(host(mk), host(0.5 * (model.metric_at(res.x + smpl)(tt) + model.metric_at(res.x - smpl)(tt)))) <= 1e-12
    # geoVI update of the residual sample
    same, st0 = nb.nonlinearly_update_residual(model, res.x, smpl, wd, wp, minimize_kwargs=dict(maxiter=0))
    assert rel(host(same), host(smpl)) <= 1e-14 and st0.status == 0
    from niftyb200 import evi
    ms, _ = nb.draw_linear_residual(M, wd, wp, from_inverse=False)
    f0, _ = evi._residual_vg(model, res.x, model.transformation(res.x), ms, res.x + smpl)
    new, st = nb.nonlinearly_update_residual(model, res.x, smpl, wd, wp, minimize_kwargs=dict(xtol=1e-10, maxiter=25,
                                                                                               cg_kwargs=dict(miniter=3)))
    assert st.status >= 0 and st.fun <= 1e-6 * max(f0, 1e-30) + 1e-18, (f0, st.fun, st.status)
    # the geoVI metric is symmetric
    u, v = dev(rng.normal(size=lay.size)), dev(rng.normal(size=lay.size))
    a_ = float(torch.dot(u, evi._geovi_metric(model, res.x, res.x + new, v)))
    b_ = float(torch.dot(evi._geovi_metric(model, res.x, res.x + new, u), v))
    assert abs(a_ - b_) <= 1e-9 * max(abs(a_), abs(b_))


def check_los(shape, dist, n_los=24, n_sampling=64, dt=np.float64, seed=8):
    """SamplingCartesianGridLOS forward / adjoint against the oracle (restatement of sampling_los.py:15-91 incl.
    jax's map_coordinates edge rule), adjointness, NaN for lines that leave the grid, shared start point; and the
    tomography metric A H f' R^T N^-1 R f' H A + 1 against the oracle composition, with a CG solve on it."""
    from oracle import los as olos
    tdt = torch.float64 if dt == np.float64 else torch.float32
    rng = np.random.default_rng(seed)
    nd = len(shape)
    ext = np.array(shape) * np.broadcast_to(np.asarray(dist, float), (nd,))
    start = rng.uniform(0.05, 0.9, size=(n_los, nd)) * ext
    end = rng.uniform(0.05, 0.9, size=(n_los, nd)) * ext
    x = rng.normal(size=shape).astype(dt)
    for st in (start, start[0]):                                  # per-line start points / one shared point (:84-90)
        R = nb.SamplingCartesianGridLOS(st, end, shape=shape, distances=dist, n_sampling_points=n_sampling, dtype=tdt)
        Ro = olos.SamplingCartesianGridLOS(st, end, shape=shape, distances=dist, n_sampling_points=n_sampling)
        assert rel(host(R(dev(x))), Ro(x.astype(np.float64))) <= TOL[dt]
        yb = rng.normal(size=n_los).astype(dt)
        assert rel(host(R.adjoint(dev(yb))), Ro.adjoint(yb.astype(np.float64))) <= TOL[dt]
        lhs = np.vdot(host(R(dev(x))).astype(np.float64), yb.astype(np.float64))
        rhs = np.vdot(x.astype(np.float64), host(R.adjoint(dev(yb))).astype(np.float64))
        assert abs(lhs - rhs) <= (1e-12 if dt == np.float64 else 1e-4) * max(abs(lhs), abs(rhs))
    # a line that leaves the grid: NaN, exactly where the reference's rule gives NaN
    bad_end = end.copy()
    bad_end[0] = ext * 1.2
    Rb = nb.SamplingCartesianGridLOS(start, bad_end, shape=shape, distances=dist, n_sampling_points=n_sampling, dtype=tdt)
    Rbo = olos.SamplingCartesianGridLOS(start, bad_end, shape=shape, distances=dist, n_sampling_points=n_sampling)
    got, ref = host(Rb(dev(x))), Rbo(x.astype(np.float64))
    assert np.array_equal(np.isnan(got), np.isnan(ref)) and np.isnan(got[0]) and not np.isnan(got[1:]).any()
    # tomography metric
    cf, co = make_pair(shape, dist, tdt, kw=GENTLE_KW)
    p = cf.init(seed)
    amps_o, azm_o = co.normalized_amplitudes(p), co.azm(p)
    tab = cf.tables_from_amplitudes(amps_o, float(azm_o))
    fp = np.exp(0.2 * rng.normal(size=shape)).astype(dt)
    ninv = rng.uniform(0.5, 2.0, size=n_los).astype(dt)
    R = nb.SamplingCartesianGridLOS(start, end, shape=shape, distances=dist, n_sampling_points=n_sampling, dtype=tdt)
    Ro = olos.SamplingCartesianGridLOS(start, end, shape=shape, distances=dist, n_sampling_points=n_sampling)
    M = nb.LOSMetric(cf, tab, R, dev(fp), dev(ninv))
    A = azm_o * co.outer_amplitude(amps_o)
    f64, n64 = fp.astype(np.float64), ninv.astype(np.float64)

    def Mo(t):
        s = f64 * co.harmonic_transform(A * t)
        z = f64 * Ro.adjoint(n64 * Ro(s))
        return A * co.harmonic_transform(z) + t
    t = rng.normal(size=shape).astype(dt)
    q, curv = M.apply_with_curvature(dev(t))
    ref = Mo(t.astype(np.float64))
    assert rel(host(q), ref) <= TOL[dt], rel(host(q), ref)
    c_ref = np.vdot(t.astype(np.float64), host(q).astype(np.float64))
    assert abs(float(curv.cpu()) - c_ref) <= 1e-12 * abs(c_ref)
    if dt == np.float64:
        j = rng.normal(size=shape)
        kw = dict(resnorm=1e-8 * float(np.linalg.norm(j)), maxiter=300)
        r, ro = cgm._cg(M, dev(j), None, **kw), ocg.cg_re(Mo, j, None, **kw)
        assert abs(r.nit - ro.nit) <= 1 and r.info == ro.info == 0 and rel(host(r.x), ro.x) <= 1e-6


def check_cg_failure_modes():
    """Non-positive-definite operator: same error behaviour as the reference
    (re/conjugate_gradient.py:151-167,276-279; conjugate_gradient.py:86-96)."""
    import pytest
    j = dev(np.ones(8))
    neg = lambda t, out=None: -t  # noqa: E731
    with pytest.raises(ValueError):
        cgm._cg(neg, j)
    assert cgm._static_cg(neg, j).info == -1
    r = cgm._cg(neg, j, _raise_nonposdef=False)
    assert r.info == 0 and r.nit == 1
    e, status = nb.ConjugateGradient(nb.GradientNormController(iteration_limit=5))(
        nb.QuadraticEnergy(j.clone(), neg, j))
    assert status == cgm.ERROR


def check_strided_scatter(dt=np.float32):
    """nb2_strided_pass_scatter == the plain strided pass followed by a permutation of rows, outer
    items and column blocks (the receiver-natural layout of the peer-memory slab exchange)."""
    tdt = torch.float32 if dt == np.float32 else torch.float64
    rng = np.random.default_rng(11)
    for shape, which in (((4, 256, 64), 1), ((256, 4, 64), 0), ((2, 512, 64), 1), ((16, 8, 32), 1), ((8, 16, 32), 0)):
        n0, n1, n2 = shape
        M = n2 // 2
        x = dev(rng.normal(size=shape).astype(dt))
        plan = get_plan(shape, None, tdt)
        ref = torch.empty_like(x)
        plan.run_pass(which, x, ref)
        ref = ref.cpu().numpy().reshape(n0, n1, M, 2)
        out = dev(np.full(shape, np.nan, dtype=dt))
        ti = lambda a: dev(np.ascontiguousarray(a, dtype=np.int64))   # noqa: E731
        if which == 1:
            rp, op = rng.permutation(n1), rng.permutation(n0)
            plan.run_strided_scatter(1, x, out, out_rowoff=ti(rp * M), out_outer_off=ti(op * n1 * M))
            got = out.cpu().numpy().reshape(n0, n1, M, 2)
            want = np.empty_like(ref)
            want[np.ix_(op, rp)] = ref
        else:
            rp, cp = rng.permutation(n0), rng.permutation(n1)
            plan.run_strided_scatter(0, x, out, out_rowoff=ti(rp * n1 * M), out_col_off=ti(cp * M))
            got = out.cpu().numpy().reshape(n0, n1, M, 2)
            want = np.empty_like(ref)
            want[np.ix_(rp, cp)] = ref
        assert np.array_equal(got, want), (shape, which, dt)


def check_fused_direction_update(shape, dist, dt=np.float64):
    """apply_with_curvature(d, pre=(r, beta)) == axpby(d = r + beta d) followed by the plain apply:
    the updated d (stored in place), M d and the curvature (nb2_pointop.pre_add)."""
    M, Mo, j, x0 = metric_pair(shape, dist, dt)
    assert M.fused_direction_update
    rng = np.random.default_rng(5)
    d0, r = rng.normal(size=shape).astype(dt), rng.normal(size=shape).astype(dt)
    beta = 0.37
    d1, d2 = dev(d0.copy()), dev(d0.copy())      # dev() is zero-copy on the CPU emulator
    q1, c1 = M.apply_with_curvature(d1, pre=(dev(r), beta))
    q1, c1 = q1.clone(), float(c1.cpu())
    cgm = sys.modules["niftyb200.cg"]
    cgm._Vec(d2).axpby(1.0, dev(r), beta, d2)
    q2, c2 = M.apply_with_curvature(d2)
    tol = 1e-13 if dt == np.float64 else 2e-6
    want = r.astype(np.float64) + beta * d0.astype(np.float64)
    assert rel(host(d1), want) <= tol and rel(host(d1), host(d2)) <= tol
    assert rel(host(q1), host(q2)) <= 10 * tol and abs(c1 - float(c2.cpu())) <= 10 * tol * abs(c1)
    # and against the oracle
    assert rel(host(q1), Mo(want)) <= TOL[dt]


def check_new_entry_point_errors():
    """Argument / support errors of the prologue update and the scatter pass surface as NativeError
    with the library's message (the C ABI returns NB2_ERR_ARG / NB2_ERR_UNSUPPORTED)."""
    from niftyb200._lib import NativeError
    x, r = dev(np.ones((8, 16))), dev(np.ones((8, 16)))
    out = torch.empty_like(x)
    plan = get_plan((8, 16), None, x.dtype)
    with pytest.raises(NativeError, match="prologue"):         # pre_add in an epilogue
        plan.hartley(x, out=out, epi=PointOp(pre_add=r, pre_scale=1.0))
    with pytest.raises(NativeError, match="alias"):            # pre_out == out
        plan.hartley(x, out=out, pro=PointOp(pre_add=r, pre_scale=1.0, pre_out=out))
    with pytest.raises(NativeError, match="alias|needs pre_add"):   # pre_out without pre_add
        plan.hartley(x, out=out, pro=PointOp(pre_out=x))
    x1 = dev(np.ones(64))
    with pytest.raises(NativeError, match="strided pass"):     # 1-D: the first pass is the row pass
        get_plan((64,), None, x1.dtype).hartley(x1, pro=PointOp(pre_add=x1.clone(), pre_scale=1.0))
    xg = dev(np.ones((6, 10)))
    with pytest.raises(NativeError, match="strided pass"):     # generic (non power-of-two) plan
        get_plan((6, 10), None, xg.dtype).hartley(xg, pro=PointOp(pre_add=xg.clone(), pre_scale=1.0))
    tab = dev(np.zeros(8, dtype=np.int64))
    with pytest.raises(NativeError, match="invalid argument"):  # in place
        plan.run_strided_scatter(1, x, x, out_rowoff=tab)
    with pytest.raises(NativeError, match="not part of this plan"):   # (8, 16) is 2-D: no axis-0 pass
        plan.run_strided_scatter(0, x, out, out_rowoff=tab)
