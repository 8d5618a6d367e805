"""CPU-only: the kernels' index arithmetic (sequential thread emulator of the SAME kernel
sources, tests/emu/README.md) and the Python host logic, checked against the oracle and the
golden vectors of the reference.  The GPU twins of these tests are in test_gpu_parity.py."""

import numpy as np
import pytest

import parity_cases as pc

pytestmark = pytest.mark.usefixtures("emu")


def test_ducc_dispatch_golden(golden):
    pc.check_hartley_golden(golden)


def test_hartley_pow2_shapes():
    pc.check_hartley_shapes([((4,), None), ((8,), None), ((64,), None), ((2048,), None), ((8192,), None),
                             ((2, 4), None), ((16, 32), None), ((128, 128), None), ((64, 16), None),
                             ((4, 8, 16), None), ((16, 16, 16), None), ((2, 2, 4), None), ((32, 4, 64), None),
                             ((3, 8, 16), (1, 2)), ((5, 64), (1,)), ((2, 3, 8, 8, 8), (2, 3, 4))])


def test_hartley_long_axes():
    # three-step radix plans (n > 1024) on strided and contiguous axes
    pc.check_hartley_shapes([((2048, 4), None), ((4096, 8), None), ((8192, 4), None), ((4, 4096), None)],
                            dtypes=(np.float64,))
    pc.check_hartley_shapes([((16384,), None), ((4096, 4), None)], dtypes=(np.float32,))


def test_c2c_fast_path():
    # every radix plan of the contiguous-axis c2c kernel (rows of 2 .. 8192), strided passes of all three kinds,
    # batch axes, axis subsets
    pc.check_c2c_shapes([((2,), None), ((4,), None), ((8,), None), ((16,), None), ((32,), None), ((64,), None), ((128,), None),
                         ((256,), None), ((512,), None), ((1024,), None), ((2048,), None), ((4096,), None), ((8192,), None),
                         ((16, 32), None), ((8, 4, 16), None), ((3, 16, 8), (1, 2)), ((5, 64), (1,)), ((2, 3, 8, 8, 4), (2, 3, 4)),
                         ((256, 8), None), ((4, 256, 16), None), ((2048, 4), None)], expect_fast=True)
    pc.check_c2c_shapes([((32, 32), None), ((16, 16, 16), None), ((64, 2), None)], dtypes=(np.complex64,), expect_fast=True)
    # not on the fast path: non-power-of-two, leading axes, 16384-point rows
    pc.check_c2c_shapes([((10,), None), ((12, 9), None), ((8, 4), (0,)), ((6, 16), None)], expect_fast=False)


def test_hartley_generic_shapes():
    pc.check_hartley_shapes([((10,), None), ((12, 9), None), ((15, 27), None), ((7, 1, 5), None), ((1,), None),
                             ((6, 16, 8), (1,)), ((8, 4, 16), (0, 2)), ((16, 2), None), ((8, 6), None)])


@pytest.mark.parametrize("shape", [(64,), (32, 16), (8, 16, 32)])
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


@pytest.mark.parametrize("shape,dist,kind", [((64, 64), 1 / 64, "power"), ((16, 16, 16), 1 / 16, "amplitude"), ((2048,), 0.01, "power")])
def test_cf_latent_vjp_directional(shape, dist, kind):
    pc.check_cf_latent_vjp_directional(shape, dist, kind)


@pytest.mark.parametrize("shape,dist,mode", [((64,), 0.3, 1), ((32, 32), 1 / 32, 1), ((16, 16, 16), 1 / 16, 1),
                                             ((8, 4, 16), 1 / 16, 2), ((5, 6), (0.3, 0.5), 2)])
def test_cf_vs_oracle(shape, dist, mode):
    pc.check_cf_vs_oracle(shape, dist, np.float64, mode)
    pc.check_cf_vs_oracle(shape, dist, np.float32, mode)


@pytest.mark.parametrize("shape,dist", [((32, 32), 1 / 32), ((8, 8, 16), (1 / 8, 1 / 8, 1 / 16))])
def test_metric(shape, dist):
    pc.check_metric(shape, dist)
    pc.check_metric(shape, dist, np.float32)


@pytest.mark.parametrize("shape,dist", [((32, 32), 1 / 32), ((16, 16, 16), 1 / 16)])
def test_cg_parity(shape, dist):
    pc.check_cg_parity(shape, dist)


def test_cg_count_identical():
    # stopping rules with separated thresholds: identical counts, both precisions, both likelihoods
    n = pc.check_cg_strict((32, 32), (1 / 32, 1 / 32), np.float64, "gaussian", min_checks=2)
    n += pc.check_cg_strict((32, 32), (1 / 32, 1 / 32), np.float32, "gaussian", min_checks=0)
    n += pc.check_cg_strict((64, 64), (1 / 64, 1 / 64), np.float64, "poisson", min_checks=6)
    n += pc.check_cg_strict((16, 16, 16), (1 / 16,) * 3, np.float32, "poisson", min_checks=4)
    assert n >= 16


def test_metric_poissonian():
    pc.check_metric((32, 32), (1 / 32, 1 / 32), np.float64, "poisson")
    pc.check_metric((8, 8, 16), (1 / 8, 1 / 8, 1 / 16), np.float32, "poisson")


@pytest.mark.parametrize("shape,dist", [((32, 32), 1 / 32), ((8, 8, 16), (1 / 8, 1 / 8, 1 / 16))])
def test_static_cg_device_resident(shape, dist):
    pc.check_static_cg_device_resident(shape, dist)
    pc.check_static_cg_device_resident(shape, dist, np.float32, "poisson" if len(shape) == 2 else "gaussian")


@pytest.mark.parametrize("shape,dist", [((32, 32), 1 / 32), ((8, 8, 16), (1 / 8, 1 / 8, 1 / 16))])
def test_sampled_metric(shape, dist):
    pc.check_sampled_metric(shape, dist)
    pc.check_sampled_metric(shape, dist, np.float32, 4, "poisson" if len(shape) == 2 else "gaussian")


def test_newton_cg_parity():
    pc.check_newton_cg_parity()


@pytest.mark.parametrize("shape,dist", [((16, 16), 1 / 16)])
def test_gaussian_model_and_geovi(shape, dist):
    pc.check_gaussian_model_and_geovi(shape, dist)


@pytest.mark.parametrize("shape,dist", [((16, 16, 16), 1 / 16), ((32, 16), (0.1, 0.3)), ((64,), 0.5)])
def test_los_response_and_tomography_metric(shape, dist):
    pc.check_los(shape, dist)
    pc.check_los(shape, dist, n_los=7, n_sampling=33, dt=np.float32)


def test_cg_failure_modes():
    pc.check_cg_failure_modes()


def test_tma_strided_pass(emu, monkeypatch):
    # persistent TMA-fed strided pass (nb2_tma.cuh): 256..1024 rows, 128-byte tiles, both precisions,
    # with the fused amplitude prologues; NB2_TMA=0 must give the same answers through StridedPass
    n0 = emu.nb2_tma_launch_count()
    shapes = [((256, 256), None), ((512, 64), None), ((1024, 128), None), ((4, 256, 64), None), ((256, 4, 64), None),
              ((256, 256, 32), None)]
    pc.check_hartley_shapes(shapes)
    assert emu.nb2_tma_launch_count() > n0
    n0 = emu.nb2_tma_launch_count()
    pc.check_cf_vs_oracle((256, 256), 1 / 256, np.float64, None)
    pc.check_cf_vs_oracle((256, 8, 64), (1 / 256, 1 / 8, 1 / 64), np.float32, None)
    pc.check_metric((256, 256), 1 / 256)
    pc.check_metric((512, 8, 32), (1 / 512, 1 / 8, 1 / 32), np.float32)
    pc.check_inplace_and_roundtrip((256, 256))
    assert emu.nb2_tma_launch_count() > n0
    n0 = emu.nb2_tma_launch_count()
    monkeypatch.setenv("NB2_TMA", "0")
    pc.clear_plans()
    pc.check_hartley_shapes(shapes[:2])
    assert emu.nb2_tma_launch_count() == n0


def test_strided_scatter(emu, monkeypatch):
    pc.check_strided_scatter(np.float32)            # default: StridedPass with the scatter tables
    monkeypatch.setenv("NB2_TMA_SCATTER", "1")
    pc.clear_plans()
    n0 = emu.nb2_tma_launch_count()
    pc.check_strided_scatter(np.float32)
    pc.check_strided_scatter(np.float64)
    assert emu.nb2_tma_launch_count() > n0          # the 256/512-row cases go through the TMA kernel
    monkeypatch.setenv("NB2_TMA", "0")
    pc.clear_plans()
    pc.check_strided_scatter(np.float32)


@pytest.mark.parametrize("shape,dist", [((32, 32), 1 / 32), ((8, 8, 16), (1 / 8, 1 / 8, 1 / 16)), ((256, 256), 1 / 256),
                                        ((512, 8, 32), (1 / 512, 1 / 8, 1 / 32))])
def test_fused_direction_update(shape, dist):
    # CG direction update d = r + beta d carried by the first pass (StridedPass on the small grids,
    # the TMA kernel with the field in shared memory on the 256/512-row ones)
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


def test_hartley_random_pow2_shapes():
    # seeded sweep over power-of-two shapes and axis subsets: every radix plan up to 2048-scalar rows, all
    # three row-pass parts, strided passes with and without a leading batch
    rng = np.random.default_rng(2024)
    cases = []
    while len(cases) < 40:
        nd = int(rng.integers(1, 4))
        shape = tuple(int(2 ** rng.integers(0, 7)) for _ in range(nd - 1)) + (int(2 ** rng.integers(1, 12)),)
        if np.prod(shape) > 1 << 16:
            continue
        first = int(rng.integers(0, nd))                    # transform the trailing axes first..nd-1
        cases.append((shape, None if first == 0 else tuple(range(first, nd))))
    pc.check_hartley_shapes(cases, dtypes=(np.float64,))
    pc.check_hartley_shapes(cases[::3], dtypes=(np.float32,))


def test_new_entry_point_errors():
    pc.check_new_entry_point_errors()


def test_hartley_leading_axes_inner_batch():
    # product-space sub-transforms: leading axes, contiguous inner batch (strided passes + mirror combine)
    import niftyb200 as nb
    import torch
    shapes = [((16, 8, 4), (0, 1)), ((8, 16, 2), (0, 1)), ((4, 32, 8), (1,)), ((32, 6), None), ((3, 8, 16, 4), (1, 2)),
              ((64, 4), (0,)), ((2, 2, 2), (0, 1)), ((16, 16, 32), (0,))]
    for shp, ax in shapes:
        if ax is not None:
            assert nb.Plan(shp, ax, torch.float64).kind == 2, (shp, ax)
    pc.check_hartley_shapes([s for s in shapes if s[1] is not None])


def test_product_hartley_joint_plan():
    # (leading axes) x (last axis) product of spaces from ONE three-pass plan (NB2_HARTLEY_PRODUCT_LAST)
    pc.check_product_hartley([(8, 16), (4, 8, 16), (16, 4, 4), (2, 2, 4), (8, 4), (4, 4, 64), (1, 4, 8), (32,),
                              (2, 4, 512)])


def test_product_correlated_field_joint_plan():
    pc.check_product_cf([((8, 16), (1 / 8, 1 / 16)), ((8,), 0.7)])
    pc.check_product_cf([((16,), 1 / 16), ((32,), 0.3)], np.float32)


def test_spectrum_cotangent_is_order_independent():
    pc.check_vjp_reproducible((16, 32), (1 / 16, 1 / 32))
    pc.check_vjp_reproducible((8, 8, 16), (1 / 8, 1 / 8, 1 / 16), np.float32)


def test_non_power_of_two_axes_bluestein():
    # generic path: axes of 32..4096 points go through the chirp-z convolution on the power-of-two row kernel,
    # shorter and longer ones through the direct sum
    pc.check_hartley_shapes([((100,), None), ((33,), None), ((96, 40), None), ((50, 36, 10), None), ((3, 48), (1,)),
                             ((45, 6), (0,)), ((1000,), None), ((4097,), None), ((4000,), None)])
    pc.check_c2c_shapes([((100,), None), ((12, 90), None), ((35, 8), (0,)), ((63, 65), None)], expect_fast=False)
    # power-of-two axes of generic plans (4-D grids, leading axes): one pass of the row kernel per axis
    pc.check_hartley_shapes([((32, 6), None), ((4, 4, 32, 64), None), ((64, 3), (0,)), ((100, 64), None),
                             ((2, 2, 2, 2, 32), None)])
    pc.check_c2c_shapes([((64, 3), (0,)), ((4, 4, 32, 64), None), ((128, 5), None)], expect_fast=False)
    # lines longer than the row kernel takes (four-step), as power-of-two axes and as padded chirp-z lines
    pc.check_hartley_shapes([((2, 32768), None), ((20000,), None), ((3, 16384, 2), (1,)), ((5000, 3), (0,))])
    pc.check_c2c_shapes([((65536,), None), ((9000,), None)], expect_fast=None)


def test_generic_path_random_shapes():
    # seeded sweep over arbitrary extents and axis subsets: direct sums, chirp-z axes, power-of-two axes of generic
    # plans, tiled and plain gathers in any combination
    rng = np.random.default_rng(77)
    cases = []
    while len(cases) < 30:
        nd = int(rng.integers(1, 4))
        shape = tuple(int(rng.integers(1, 70)) if rng.random() < 0.6 else int(rng.integers(70, 300)) for _ in range(nd))
        if np.prod(shape) > 40000 or np.prod(shape) < 2:
            continue
        k = int(rng.integers(1, nd + 1))
        cases.append((shape, tuple(sorted(rng.choice(nd, size=k, replace=False).tolist()))))
    pc.check_hartley_shapes(cases)
    pc.check_c2c_shapes(cases[:12], expect_fast=None)


def test_matern_latent_derivatives():
    pc.check_matern_latent_derivatives((16, 16), (1 / 16, 1 / 16))
    pc.check_matern_latent_derivatives((8, 8, 8), (1 / 8,) * 3, kind="power", renorm=False)
    pc.check_matern_latent_derivatives((24,), (0.1,), kind="power")
