"""Pin the CPU oracle against golden vectors generated from the real reference
(``tests/golden/make_golden.py``).  CPU-only; no CUDA involved."""

import numpy as np
import pytest
from numpy.testing import assert_allclose

import oracle
from oracle import cg as ocg
from oracle import correlated_field as ocf
from oracle import geometry as ogeo
from oracle import transforms as otr

CONVS = ("non_canonical_hartley", "canonical_hartley")


def _axes(a):
    a = tuple(int(x) for x in a)
    return None if a == (-1,) else a


def test_hartley_fft_golden(golden):
    g = golden["transforms"]
    for n in range(int(g["ncases"])):
        a, axes = g[f"h{n}_in"], _axes(g[f"h{n}_axes"])
        for c in CONVS:
            assert_allclose(otr.hartley(a, axes, c), g[f"h{n}_{c}"], rtol=1e-13, atol=1e-13)
        if f"h{n}_cin" in g:
            z = g[f"h{n}_cin"]
            assert_allclose(otr.fftn(z, axes), g[f"h{n}_fftn"], rtol=1e-13, atol=1e-13)
            assert_allclose(otr.ifftn(z, axes), g[f"h{n}_ifftn"], rtol=1e-13, atol=1e-13)


def test_direct_hartley_samples(golden):
    g = golden["transforms"]
    a = g["h5_in"]  # (8, 6, 10)
    idx = [(0, 0, 0), (3, 5, 9), (7, 1, 4)]
    for c in CONVS:
        ref = g[f"h5_{c}"]
        got = otr.direct_hartley_samples(a, idx, c)
        assert_allclose(got, [ref[i] for i in idx], rtol=1e-12, atol=1e-12)


def test_hartley_operator_golden(golden):
    g = golden["operator"]
    for n in range(int(g["ncases"])):
        x, dist = g[f"o{n}_in"], g[f"o{n}_dist"]
        shape = x.shape
        dvol = float(np.prod(dist))
        hvol = float(np.prod(1.0 / (np.array(shape) * dist)))
        axes = tuple(range(x.ndim))
        y = otr.hartley_operator(x, axes, dvol)
        assert_allclose(y, g[f"o{n}_times"], rtol=1e-13, atol=1e-13)
        assert_allclose(otr.hartley_operator(y, axes, hvol), g[f"o{n}_inverse"], rtol=1e-12, atol=1e-12)
        assert_allclose(otr.hartley_operator(x, axes, dvol), g[f"o{n}_adjoint"], rtol=1e-13, atol=1e-13)
        assert_allclose(otr.hartley_operator(g[f"o{n}_cin"], axes, dvol), g[f"o{n}_ctimes"],
                        rtol=1e-13, atol=1e-13)
        assert_allclose(otr.fftn(g[f"o{n}_cin"]) * dvol, g[f"o{n}_fft_times"], rtol=1e-13, atol=1e-13)


# the CF cases of make_golden.CF_CASES, restated for the oracle's (re-style) maker
CF_CASES = {
    "cf1d": ([((16,), 0.1, dict(fluctuations=(1., .1), loglogavgslope=(-1., .1),
                                flexibility=(1., .1), asperity=(.2, 2e-2)))], 0.0, (.1, .1)),
    "cf2d": ([((12, 12), 1. / 12, dict(fluctuations=(1e-1, 5e-3), loglogavgslope=(-1., 1e-2),
                                       flexibility=(1., .5), asperity=(.5, 5e-2)))], 0.0, (1e-3, 1e-4)),
    "cf2d_odd": ([((5, 6), (0.3, 0.5), dict(fluctuations=(3., 2.), loglogavgslope=(-2., 1.),
                                            flexibility=(3., 2.), asperity=None))], 1.5, (.1, .1)),
    "cf3d": ([((8, 8, 8), 1. / 8, dict(fluctuations=(1., .5), loglogavgslope=(-5., .2),
                                       flexibility=(1., .2), asperity=(.5, 5e-2)))], 2.0, (1e-1, 3e-2)),
    "cfprod": ([((8, 8), 1. / 8, dict(fluctuations=(1e-1, 5e-3), loglogavgslope=(-1., 1e-2),
                                      flexibility=(1., .5), asperity=(.5, 5e-2))),
                ((6,), 0.7, dict(fluctuations=(1., 1e-2), loglogavgslope=(-4., 1.),
                                 flexibility=(3.1, 1.), asperity=(.5, .1)))], 0.5, (1e-2, 1e-3)),
}


def build_oracle_cf(name):
    subs, off_mean, off_std = CF_CASES[name]
    mk = ocf.CorrelatedFieldMaker("")
    mk.set_amplitude_total_offset(off_mean, off_std)
    for i, (shape, dist, kw) in enumerate(subs):
        pre = "" if len(subs) == 1 else f"ax{i}"
        # numpy stack parametrises the *power* spectrum (test_re/test_correlated_field.py:150)
        mk.add_fluctuations(shape, dist, **kw, prefix=pre, non_parametric_kind="power")
    return mk.finalize()


def golden_pos(g, tag, cf, kind="pos"):
    p = {}
    for k in cf.domain:
        v = g[f"{tag}_{kind}_{k}"]
        p[k] = v.T if k.endswith("spectrum") else v  # test_re/test_correlated_field.py:163
    return p


@pytest.mark.parametrize("name", sorted(CF_CASES))
@pytest.mark.parametrize("conv", CONVS)
def test_cf_forward_and_tables_golden(golden, name, conv):
    g = golden["correlated_field"]
    tag = f"{name}_{conv}"
    cf = build_oracle_cf(name)
    p = golden_pos(g, tag, cf)
    assert_allclose(cf(p, conv), g[f"{tag}_val"], rtol=1e-10, atol=1e-12)
    amps = cf.normalized_amplitudes(p)
    assert_allclose(cf.azm(p), g[f"{tag}_azm"], rtol=1e-13)
    for i, (a, grid) in enumerate(zip(amps, cf.grids)):
        assert np.array_equal(grid.power_distributor, g[f"{tag}_pindex{i}"])
        # entry 0 differs by design: the numpy stack leaves a junk zero-mode amplitude that
        # is later masked (correlated_fields.py:191-201), re sets totvol (:303); the
        # product azm*amp[0] is what enters the field and is checked through `val`.
        assert_allclose(a[1:], g[f"{tag}_namp{i}"][1:], rtol=1e-10)
    # hot tail with fixed tables == full model
    assert_allclose(cf.apply(amps, cf.azm(p), p["xi"], conv), g[f"{tag}_val"], rtol=1e-10, atol=1e-12)


@pytest.mark.parametrize("name", sorted(CF_CASES))
@pytest.mark.parametrize("conv", CONVS)
def test_cf_adjoint_xi_golden(golden, name, conv):
    """xi-block of the reference Jacobian adjoint == oracle adjoint (fixed tables)."""
    g = golden["correlated_field"]
    tag = f"{name}_{conv}"
    cf = build_oracle_cf(name)
    p = golden_pos(g, tag, cf)
    amps, azm = cf.normalized_amplitudes(p), cf.azm(p)
    xibar, ampbars, azmbar, offbar = cf.adjoint(amps, azm, g[f"{tag}_ybar"], p["xi"], conv)
    assert_allclose(xibar, g[f"{tag}_adj_xi"], rtol=1e-10, atol=1e-12)
    # adjointness of the table cotangents: <ybar, d apply/d amp . da> == <ampbar, da>
    rng = np.random.default_rng(3)
    for i in range(len(amps)):
        da = rng.normal(size=amps[i].shape)
        amps2 = [a.copy() for a in amps]
        eps = 1e-6
        amps2[i] = amps[i] + eps * da
        fd = (cf.apply(amps2, azm, p["xi"], conv) - cf.apply(amps, azm, p["xi"], conv)) / eps
        assert_allclose(np.vdot(g[f"{tag}_ybar"], fd), np.vdot(ampbars[i], da), rtol=1e-5)
    fd = (cf.apply(amps, azm * (1 + 1e-6), p["xi"], conv) - cf.apply(amps, azm, p["xi"], conv)) / (azm * 1e-6)
    assert_allclose(np.vdot(g[f"{tag}_ybar"], fd), azmbar, rtol=1e-5)
    assert_allclose(offbar, g[f"{tag}_ybar"].sum(), rtol=1e-13)


def test_power_index_np_matches_re_rule():
    for shape, dist in [((16,), 0.1), ((12, 12), 1 / 12), ((5, 6), (0.3, 0.5)), ((8, 8, 8), 1 / 8),
                        ((6, 10, 4), (1.0, 0.5, 0.25))]:
        pi = ogeo.power_index_np(shape, dist)
        idx, um, cnt = ogeo.fourier_mode_distributor(shape, dist)
        assert np.array_equal(pi.pindex, idx)
        assert np.array_equal(pi.rho, cnt)
        assert_allclose(pi.unique_k, um, rtol=1e-12)


def _cg_case(g, n):
    A, w, x0, b, dist = (g[f"cg{n}_{k}"] for k in ("A", "w", "x0", "b", "dist"))
    shape = A.shape
    # harmonic scalar_dvol exactly as RGSpace computes it (rg_space.py:77-82)
    fct = float(np.prod(1.0 / (np.array(shape) * float(dist))))
    axes = tuple(range(A.ndim))

    def H(x):
        return otr.hartley(x, axes) * fct

    def mat(t):
        return A * H(w * H(A * t)) + t
    return mat, x0, b


def test_cg_numpy_stack_golden(golden):
    g = golden["cg"]
    for n in range(int(g["ncases"])):
        mat, x0, b = _cg_case(g, n)
        assert_allclose(mat(x0), g[f"cg{n}_Mx0"], rtol=1e-11, atol=1e-12)
        ctrls = dict(absdelta=ocg.AbsDeltaEnergyController(1e-6 * x0.size, iteration_limit=100),
                     gradnorm=ocg.GradientNormController(tol_rel_gradnorm=1e-4, iteration_limit=100),
                     limit7=ocg.AbsDeltaEnergyController(0., iteration_limit=7),
                     reset=ocg.GradientNormController(tol_abs_gradnorm=1e-9, iteration_limit=45))
        for cname, ctrl in ctrls.items():
            en = ocg.QuadraticEnergy(x0, mat, b)
            res, status, nit = ocg.conjugate_gradient_np(en, ctrl)
            tag = f"cg{n}_{cname}"
            assert status == int(g[f"{tag}_status"])
            assert ctrl._itcount == int(g[f"{tag}_nit"]), (tag, ctrl._itcount, int(g[f"{tag}_nit"]))
            assert_allclose(res.position, g[f"{tag}_x"], rtol=1e-8, atol=1e-10)
            assert_allclose(res.value, float(g[f"{tag}_energy"]), rtol=1e-10)


def test_cg_re_variants_agree(golden):
    """``_cg`` == ``_static_cg`` (the reference's ``test/test_re/test_ncg.py:109-187`` idea) and
    both solve the system; jax is absent so this pins the two restatements to each other
    and to the numpy-stack golden solution."""
    g = golden["cg"]
    for n in range(int(g["ncases"])):
        mat, x0, b = _cg_case(g, n)
        r1 = ocg.cg_re(mat, b, x0, absdelta=1e-9, maxiter=200)
        r2 = ocg.static_cg_re(mat, b, x0, absdelta=1e-9, maxiter=200)
        assert r1.nit == r2.nit and r1.info == r2.info == 0 and r1.nfev == r2.nfev
        assert_allclose(r1.x, r2.x, rtol=1e-12, atol=1e-13)
        assert_allclose(mat(r1.x), b, rtol=5e-3, atol=5e-3)
        r3 = ocg.cg_re(mat, b, None, resnorm=1e-8 * np.linalg.norm(b), maxiter=7)
        r4 = ocg.static_cg_re(mat, b, None, resnorm=1e-8 * np.linalg.norm(b), maxiter=7)
        assert r3.nit == r4.nit == 7 and r3.info == r4.info == 7


def test_cg_nonposdef():
    mat = lambda t: -t  # noqa: E731
    j = np.ones(5)
    with pytest.raises(ValueError):
        ocg.cg_re(mat, j)
    r = ocg.static_cg_re(mat, j)
    assert r.info == -1
    assert oracle is not None


def test_isotropic_shortcut_equals_restated_algorithm():
    """oracle.geometry uses an integer-ksq shortcut above 2^24 points (where the reference's float sort takes
    minutes): same index, same mode lengths (bit for bit), same multiplicities as the restated
    ``get_fourier_mode_distributor`` on every isotropic shape tried."""
    from oracle import geometry as g
    for shape, d in (((64, 64), 1 / 64), ((32, 32, 32), 1 / 32), ((16, 64), (1 / 16, 1 / 64)), ((128, 256), (1 / 128, 1 / 256)),
                     ((8, 16, 32), (1 / 8, 1 / 16, 1 / 32)), ((12, 12), 0.3)):
        idx, um, cnt = g.fourier_mode_distributor(shape, d)
        idx2, um2, cnt2 = g.fourier_mode_distributor_isotropic(shape, d)
        assert np.array_equal(idx, idx2) and np.array_equal(cnt, cnt2)
        assert np.array_equal(um2, um)          # bit for bit: log-differences of these amplify one ulp by 1e9


def test_direct_hartley_samples_equal_the_full_transform():
    from oracle import transforms as otr
    rng = np.random.default_rng(0)
    for shape in ((37,), (12, 9), (8, 6, 10), (4, 5, 6, 7)):
        a = rng.normal(size=shape)
        idx = [tuple(int(rng.integers(0, n)) for n in shape) for _ in range(5)]
        for c in (otr.NON_CANONICAL, otr.CANONICAL):
            ref = otr.hartley(a, None, c)
            got = otr.direct_hartley_samples(a, idx, c)
            assert np.max(np.abs(got - np.array([ref[i] for i in idx]))) <= 1e-12 * np.sqrt(a.size) * 10
