"""GPU parity tests proper: the CUDA kernels, called through the C ABI of
``include/niftyb200.h`` (ctypes), against the CPU oracle and the reference's golden vectors,
plus size-independent properties at the BASELINE sizes (sub-sampled direct Hartley sums,
Parseval, involution, adjointness)."""

import numpy as np
import pytest
import torch

import parity_cases as pc

pytestmark = [pytest.mark.gpu, pytest.mark.usefixtures("cuda")]


def test_native_library_loaded(cuda):
    import ctypes
    assert cuda.path.endswith("libniftyb200.so") and cuda.nb2_version() >= 100
    assert cuda.nb2_sm_count() > 0
    assert isinstance(cuda.cdll, ctypes.CDLL)


def test_ducc_dispatch_golden(golden):
    pc.check_hartley_golden(golden)


def test_hartley_pow2_shapes():
    pc.check_hartley_shapes([((4,), None), ((8,), None), ((64,), None), ((2048,), None), ((8192,), None),
                             ((2, 4), None), ((16, 32), None), ((128, 128), None), ((64, 16), None),
                             ((4, 8, 16), None), ((16, 16, 16), None), ((2, 2, 4), None), ((32, 4, 64), None),
                             ((3, 8, 16), (1, 2)), ((5, 64), (1,)), ((2, 3, 8, 8, 8), (2, 3, 4)),
                             ((1024, 1024), None), ((64, 64, 64), None), ((128, 256, 64), None),
                             ((512, 8, 1024), None), ((8, 1024, 64), None)])


def test_hartley_long_axes():
    pc.check_hartley_shapes([((2048, 64), None), ((4096, 64), None), ((8192, 16), None), ((16, 4096), None),
                             ((64, 8192), None), ((16384,), None), ((2048, 2048), None)])


def test_c2c_fast_path():
    # fftn / ifftn on the strided passes + the contiguous-axis c2c kernel: every radix plan, TMA-fed strided
    # passes (256..1024 rows), three-step plans, batch axes, axis subsets
    pc.check_c2c_shapes([((2,), None), ((4,), None), ((8,), None), ((16,), None), ((32,), None), ((64,), None), ((128,), None),
                         ((256,), None), ((512,), None), ((1024,), None), ((2048,), None), ((4096,), None), ((8192,), None),
                         ((16, 32), None), ((8, 4, 16), None), ((3, 16, 8), (1, 2)), ((5, 64), (1,)), ((2, 3, 8, 8, 4), (2, 3, 4)),
                         ((1024, 1024), None), ((64, 4096), None), ((4096, 64), None), ((512, 8, 1024), None),
                         ((8, 1024, 64), None), ((128, 128, 128), None)], expect_fast=True)
    pc.check_c2c_shapes([((256, 256, 256), None), ((2048, 2048), None)], dtypes=(np.complex64,), expect_fast=True)
    pc.check_c2c_shapes([((10,), None), ((12, 9), None), ((8, 4), (0,)), ((6, 16), None)], expect_fast=False)


def test_hartley_generic_shapes():
    pc.check_hartley_shapes([((10,), None), ((12, 9), None), ((15, 27), None), ((7, 1, 5), None), ((1,), None),
                             ((6, 16, 8), (1,)), ((8, 4, 16), (0, 2)), ((16, 2), None), ((8, 6), None),
                             ((100, 60), None)])


@pytest.mark.parametrize("shape", [(64,), (32, 16), (8, 16, 32), (256, 256), (128, 128, 128)])
def test_inplace_roundtrip(shape):
    pc.check_inplace_and_roundtrip(shape)
    pc.check_inplace_and_roundtrip(shape, np.float32)


def test_operator_golden(golden):
    pc.check_operator_golden(golden)


def test_vdot():
    pc.check_vdot()


def test_cf_golden(golden):
    pc.check_cf_golden(golden, ["cf1d", "cf2d", "cf2d_odd", "cf3d", "cfprod"])


def test_cf_latent_vjp_golden(golden):
    pc.check_cf_latent_vjp_golden(golden, ["cf1d", "cf2d", "cf2d_odd", "cf3d", "cfprod"])


def test_latent_metric_golden(golden):
    pc.check_latent_metric_golden(golden, ["cf1d", "cf2d", "cf2d_odd", "cf3d", "cfprod"])


@pytest.mark.parametrize("shape,dist,kind", [((64, 64), 1 / 64, "power"), ((16, 16, 16), 1 / 16, "amplitude"), ((2048,), 0.01, "power"), ((512, 512), 1 / 512, "power"), ((64, 64, 64), 1 / 64, "amplitude")])
def test_cf_latent_vjp_directional(shape, dist, kind):
    pc.check_cf_latent_vjp_directional(shape, dist, kind)


@pytest.mark.parametrize("shape,dist,mode", [((64,), 0.3, 1), ((32, 32), 1 / 32, 1), ((16, 16, 16), 1 / 16, 1),
                                             ((8, 4, 16), 1 / 16, 2), ((5, 6), (0.3, 0.5), 2),
                                             ((128, 128), 1 / 128, 1), ((512, 512), 1 / 512, 1),
                                             ((64, 64, 64), 1 / 64, 1), ((128, 128, 128), 1 / 128, 1)])
def test_cf_vs_oracle(shape, dist, mode):
    pc.check_cf_vs_oracle(shape, dist, np.float64, mode)
    pc.check_cf_vs_oracle(shape, dist, np.float32, mode)


@pytest.mark.parametrize("shape,dist", [((32, 32), 1 / 32), ((8, 8, 16), (1 / 8, 1 / 8, 1 / 16)),
                                        ((128, 128), 1 / 128), ((64, 64, 64), 1 / 64)])
def test_metric(shape, dist):
    pc.check_metric(shape, dist)
    pc.check_metric(shape, dist, np.float32)


@pytest.mark.parametrize("shape,dist", [((32, 32), 1 / 32), ((16, 16, 16), 1 / 16), ((128, 128), 1 / 128),
                                        ((64, 64, 64), 1 / 64)])
def test_cg_parity(shape, dist):
    pc.check_cg_parity(shape, dist, strict=int(np.prod(shape)) <= 4096)


@pytest.mark.parametrize("shape", [(128, 128), (64, 64, 64), (512, 512), (128, 128, 128)])
def test_cg_count_identical(shape):
    """IDENTICAL nit / info / nfev for `_cg`, `_static_cg` and `ConjugateGradient` on stopping rules with
    separated thresholds (parity_cases.check_cg_strict), fp64 and fp32, Gaussian (tomography-style) and
    Poissonian (C2-style, w = exp(s)) metrics."""
    dist = tuple(1.0 / s for s in shape)
    big = int(np.prod(shape)) > 2 ** 20
    n = pc.check_cg_strict(shape, dist, np.float64, "poisson", min_checks=8)
    n += pc.check_cg_strict(shape, dist, np.float32, "poisson", min_checks=6)
    if not big:
        n += pc.check_cg_strict(shape, dist, np.float64, "gaussian", min_checks=2)
        n += pc.check_cg_strict(shape, dist, np.float32, "gaussian", min_checks=0)
    assert n >= 16


@pytest.mark.parametrize("shape", [(64, 64), (256, 256), (32, 32, 32)])
def test_metric_poissonian(shape):
    dist = tuple(1.0 / s for s in shape)
    pc.check_metric(shape, dist, np.float64, "poisson")
    pc.check_metric(shape, dist, np.float32, "poisson")


@pytest.mark.parametrize("shape,dist", [((32, 32), 1 / 32), ((128, 128), 1 / 128), ((64, 64, 64), 1 / 64), ((256, 8, 256), (1 / 256, 1 / 8, 1 / 256))])
def test_static_cg_device_resident(shape, dist):
    pc.check_static_cg_device_resident(shape, dist)
    pc.check_static_cg_device_resident(shape, dist, np.float32, "poisson" if len(shape) == 2 else "gaussian")


@pytest.mark.parametrize("shape,dist", [((32, 32), 1 / 32), ((256, 256), 1 / 256), ((64, 64, 64), 1 / 64), ((256, 16, 256), (1 / 256, 1 / 16, 1 / 256))])
def test_sampled_metric(shape, dist):
    pc.check_sampled_metric(shape, dist)
    pc.check_sampled_metric(shape, dist, np.float32, 4, "poisson" if len(shape) == 2 else "gaussian")


def test_newton_cg_parity():
    pc.check_newton_cg_parity()


@pytest.mark.parametrize("shape,dist", [((16, 16), 1 / 16), ((64, 64), 1 / 64), ((16, 16, 16), 1 / 16)])
def test_gaussian_model_and_geovi(shape, dist):
    pc.check_gaussian_model_and_geovi(shape, dist)


@pytest.mark.parametrize("shape,dist", [((16, 16, 16), 1 / 16), ((32, 16), (0.1, 0.3)), ((64,), 0.5), ((64, 64, 64), 1 / 64), ((128, 128, 128), 1 / 128)])
def test_los_response_and_tomography_metric(shape, dist):
    pc.check_los(shape, dist)
    pc.check_los(shape, dist, n_los=7, n_sampling=33, dt=np.float32)


def test_cg_failure_modes():
    pc.check_cg_failure_modes()


# ---------------------------------------------------------------- BASELINE sizes
@pytest.mark.parametrize("shape,dt", [((4096, 4096), np.float64), ((4096, 4096), np.float32),
                                      ((512, 512, 512), np.float64), ((512, 512, 512), np.float32)])
def test_baseline_size_vs_oracle(shape, dt):
    """configs[1] (4096^2) and configs[2] (512^3): the CUDA path against the full oracle."""
    pc.check_baseline_size(shape, dt)


@pytest.mark.parametrize("shape,dt", [((4096, 4096), torch.float32), ((4096, 4096), torch.float64),
                                      ((512, 512, 512), torch.float32), ((512, 512, 512), torch.float64),
                                      ((1024, 1024, 1024), torch.float32)])
def test_full_size_properties(shape, dt):
    """Sub-sampled direct sums (16 output elements, O(N) each, evaluated slab by slab on the host), Parseval
    ||Hx||^2 = N ||x||^2 and H(H x) = N x at full size."""
    from niftyb200.transforms import PointOp, get_plan
    from oracle import transforms as otr
    n = int(np.prod(shape))
    gen = torch.Generator(device="cuda").manual_seed(42)
    x = torch.randn(shape, dtype=dt, device="cuda", generator=gen)
    plan = get_plan(shape, None, dt)
    y = plan.hartley(x)
    tol = 1e-5 if dt == torch.float32 else 1e-12
    nx, ny = float(torch.linalg.vector_norm(x.double())), float(torch.linalg.vector_norm(y.double()))
    assert abs(ny ** 2 / (n * nx ** 2) - 1.0) <= 10 * tol
    rng = np.random.default_rng(1)
    idx = [tuple(0 for _ in shape), tuple(s // 2 for s in shape), tuple(s - 1 for s in shape),
           tuple((7 * (i + 1) * s) // 23 for i, s in enumerate(shape))]
    idx += [tuple(int(rng.integers(0, s)) for s in shape) for _ in range(12)]
    ref = otr.direct_hartley_samples(x.cpu().numpy(), idx)
    got = np.array([float(y[i]) for i in idx])
    # element-wise error of an N-point transform with rel-L2 accuracy tol: ~ tol * rms(y)
    assert np.max(np.abs(got - ref)) <= 30 * tol * ny / np.sqrt(n), (np.abs(got - ref), ny / np.sqrt(n))
    z = plan.hartley(y, out=y, epi=PointOp(scale=1.0 / n))
    err = float(torch.linalg.vector_norm((z - x).double())) / nx
    assert err <= 4 * tol, err


def test_headline_grid_vs_oracle():
    """configs[3] grid, 1024^3 fp32: the CUDA Hartley transform against the full oracle transform (SciPy
    pocketfft on all host cores, evaluated in fp64 when the host has the memory for it, else in fp32 --
    still an independent implementation), rel-L2 <= 1e-5.  Needs ~45 GB (fp32 oracle) / ~75 GB (fp64) of host RAM."""
    import os
    import psutil
    import scipy.fft
    avail = psutil.virtual_memory().available
    if avail < 48 << 30:
        pytest.skip(f"host has {avail >> 30} GB available; the 1024^3 oracle transform needs 48 GB")
    from niftyb200.transforms import get_plan
    shape = (1024, 1024, 1024)
    gen = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(shape, dtype=torch.float32, device="cuda", generator=gen)
    y = get_plan(shape, None, torch.float32).hartley(x)
    xh = x.cpu().numpy()
    del x
    yh = y.cpu().numpy()
    del y
    torch.cuda.empty_cache()
    f = scipy.fft.fftn(xh.astype(np.float64) if avail >= 96 << 30 else xh, workers=os.cpu_count())
    del xh
    num = den = 0.0
    for k in range(shape[0]):                 # plane by plane: no further full-size temporaries
        ref = f[k].real.astype(np.float64) + f[k].imag
        d = yh[k].astype(np.float64) - ref
        num += float(np.vdot(d, d))
        den += float(np.vdot(ref, ref))
    assert np.sqrt(num / den) <= 1e-5, np.sqrt(num / den)


def test_tma_strided_pass(cuda, monkeypatch):
    # persistent TMA-fed strided pass (nb2_tma.cuh): 256..1024 rows, both precisions, fused prologues;
    # NB2_TMA=0 must give the same answers through StridedPass
    n0 = cuda.nb2_tma_launch_count()
    shapes = [((1024, 16, 1024), None), ((16, 1024, 1024), None), ((512, 32, 512), None), ((32, 512, 512), None),
              ((256, 64, 512), None), ((64, 256, 512), None), ((1024, 16384), None), ((256, 256, 256), None)]
    pc.check_hartley_shapes(shapes)
    assert cuda.nb2_tma_launch_count() >= n0 + 2 * len(shapes)
    n0 = cuda.nb2_tma_launch_count()
    pc.check_cf_vs_oracle((256, 64, 512), (1 / 256, 1 / 64, 1 / 512), np.float64, None)
    pc.check_cf_vs_oracle((1024, 16, 1024), (1 / 1024, 1 / 16, 1 / 1024), np.float32, None)
    pc.check_metric((512, 32, 512), (1 / 512, 1 / 32, 1 / 512))
    pc.check_metric((1024, 16, 1024), (1 / 1024, 1 / 16, 1 / 1024), np.float32)
    pc.check_inplace_and_roundtrip((256, 256, 256), np.float32)
    pc.check_inplace_and_roundtrip((512, 32, 512))
    assert cuda.nb2_tma_launch_count() > n0
    n0 = cuda.nb2_tma_launch_count()
    monkeypatch.setenv("NB2_TMA", "0")
    pc.clear_plans()
    pc.check_hartley_shapes(shapes[:2])
    assert cuda.nb2_tma_launch_count() == n0


def test_strided_scatter(cuda, monkeypatch):
    pc.check_strided_scatter(np.float32)
    monkeypatch.setenv("NB2_TMA_SCATTER", "1")
    pc.clear_plans()
    pc.check_strided_scatter(np.float32)
    pc.check_strided_scatter(np.float64)
    monkeypatch.setenv("NB2_TMA", "0")
    pc.clear_plans()
    pc.check_strided_scatter(np.float32)


@pytest.mark.parametrize("shape,dist", [((32, 32), 1 / 32), ((8, 8, 16), (1 / 8, 1 / 8, 1 / 16)), ((4096, 256), 1 / 4096),
                                        ((1024, 16, 1024), (1 / 1024, 1 / 16, 1 / 1024)), ((256, 256, 256), 1 / 256)])
def test_fused_direction_update(shape, dist):
    pc.check_fused_direction_update(shape, dist)
    pc.check_fused_direction_update(shape, dist, np.float32)


def test_row_pass_compile_time_epilogues():
    # rows of 512 / 1024 / 2048 scalars take the RowPass instantiations with compile-time epilogues
    # (plain, * weight, * field + add + dot)
    pc.check_hartley_shapes([((512,), None), ((4, 1024), None), ((2, 2, 2048), None)])
    pc.check_metric((8, 512), (1 / 8, 1 / 512))
    pc.check_metric((4, 4, 1024), (1 / 4, 1 / 4, 1 / 1024), np.float32)
    pc.check_cf_vs_oracle((2, 2048), (1 / 2, 1 / 2048), np.float64, None)
    pc.check_fused_direction_update((8, 512), (1 / 8, 1 / 512))


def test_new_entry_point_errors():
    pc.check_new_entry_point_errors()


def test_hartley_leading_axes_inner_batch():
    import niftyb200 as nb
    shapes = [((16, 8, 4), (0, 1)), ((64, 128, 32), (0, 1)), ((4, 32, 8), (1,)), ((3, 8, 16, 4), (1, 2)),
              ((1024, 16), (0,)), ((2, 2, 2), (0, 1)), ((256, 256, 2), (0, 1))]
    for shp, ax in shapes:
        assert nb.Plan(shp, ax, torch.float64).kind == 2, (shp, ax)
    pc.check_hartley_shapes(shapes)


@pytest.mark.parametrize("dt", [torch.float32])
def test_product_space_c5_properties(dt):
    """BASELINE configs[4] grid (2048 x 2048 space x 128 frequency): the two sub-space transforms
    (space: leading axes + inner batch; frequency: trailing axis) at full size -- Parseval and
    involution of each, and of their composition."""
    from niftyb200.transforms import PointOp, get_plan
    shape = (2048, 2048, 128)
    gen = torch.Generator(device="cuda").manual_seed(7)
    x = torch.randn(shape, dtype=dt, device="cuda", generator=gen)
    nx = float(torch.linalg.vector_norm(x.double()))
    ps, pf = get_plan(shape, (0, 1), dt), get_plan(shape, (2,), dt)
    assert ps.kind == 2 and pf.kind == 1
    tol = 1e-5
    y = ps.hartley(x)
    ny = float(torch.linalg.vector_norm(y.double()))
    assert abs(ny ** 2 / (2048 * 2048 * nx ** 2) - 1.0) <= 10 * tol
    # one frequency channel against the 2-D trailing-axes transform of that channel
    ref = get_plan((2048, 2048), None, dt).hartley(x[:, :, 5].contiguous())
    assert float(torch.linalg.vector_norm((y[:, :, 5] - ref).double()) / torch.linalg.vector_norm(ref.double())) <= tol
    y = pf.hartley(y, out=y)
    # the joint three-pass plan (NB2_HARTLEY_PRODUCT_LAST) == the two sub-space transforms in a row
    yj = get_plan(shape, None, dt).hartley(x, product_last=True)
    errj = float(torch.linalg.vector_norm((yj - y).double()) / torch.linalg.vector_norm(y.double()))
    assert errj <= 2 * tol, errj
    del yj
    y = pf.hartley(y, out=y, epi=PointOp(scale=1.0 / 128))
    y = ps.hartley(y, out=y, epi=PointOp(scale=1.0 / (2048 * 2048)))
    err = float(torch.linalg.vector_norm((y - x).double())) / nx
    assert err <= 4 * tol, err


@pytest.mark.gpu
def test_product_hartley_joint_plan():
    # (leading axes) x (last axis) product of spaces from ONE three-pass plan, incl. the TMA strided passes
    # (256..1024-point axes) and three-step rows
    pc.check_product_hartley([(8, 16), (4, 8, 16), (16, 4, 4), (2, 2, 4), (8, 4), (4, 4, 64), (1, 4, 8), (32,),
                              (2, 4, 512), (256, 256, 64), (512, 32, 128), (64, 2048, 16), (16, 8192)])


@pytest.mark.gpu
def test_product_correlated_field_joint_plan():
    pc.check_product_cf([((8, 16), (1 / 8, 1 / 16)), ((8,), 0.7)])
    pc.check_product_cf([((16,), 1 / 16), ((32,), 0.3)], np.float32)
    pc.check_product_cf([((256, 256), (1 / 256, 1 / 256)), ((64,), 0.7)], np.float32)
    pc.check_product_cf([((128, 64), (1 / 128, 1 / 64)), ((32,), 0.7)])


def test_spectrum_cotangent_is_order_independent():
    # thousands of CTAs race on the same bins: the fixed-point accumulation makes the result bit-reproducible
    pc.check_vjp_reproducible((512, 512), (1 / 512, 1 / 512))
    pc.check_vjp_reproducible((128, 128, 128), (1 / 128,) * 3, np.float32)
    pc.check_vjp_reproducible((16, 32), (1 / 16, 1 / 32))


def test_non_power_of_two_axes_bluestein():
    pc.check_hartley_shapes([((100,), None), ((33,), None), ((96, 40), None), ((50, 36, 10), None), ((3, 48), (1,)),
                             ((45, 6), (0,)), ((1000, 1000), None), ((4097,), None), ((4000, 3), None), ((384, 96, 100), None)])
    pc.check_c2c_shapes([((100,), None), ((12, 90), None), ((35, 8), (0,)), ((63, 65), None), ((1000, 300), None)],
                        expect_fast=False)
    pc.check_cf_vs_oracle((96, 100), (1 / 96, 1 / 100), np.float64, None)
    pc.check_hartley_shapes([((32, 6), None), ((16, 16, 32, 64), None), ((64, 3), (0,)), ((8192, 3), (0,)), ((100, 64), None)])
    pc.check_c2c_shapes([((64, 3), (0,)), ((4, 4, 32, 64), None), ((128, 5), None)], expect_fast=False)
    # long lines: four-step transforms of 2^15 .. 2^20 points, chirp-z lines of 2^16 .. 2^21 points
    pc.check_hartley_shapes([((2, 32768), None), ((20000,), None), ((3, 16384, 2), (1,)), ((5000, 3), (0,)),
                             ((1 << 20,), None), ((1000003,), None)])
    pc.check_c2c_shapes([((65536,), None), ((9000,), None), ((300000,), None)], expect_fast=None)


def test_matern_latent_derivatives():
    pc.check_matern_latent_derivatives((64, 64), (1 / 64, 1 / 64))
    pc.check_matern_latent_derivatives((16, 16, 16), (1 / 16,) * 3, kind="power", renorm=False)
