"""BASELINE configs[4] on one GPU: product-space correlated field 2048 x 2048 (space) x 128
(frequency), two-axis amplitudes.  Times the CF forward / adjoint apply and one metric-CG
iteration (CUDA events).  Usage: python tools/prof_c5.py [f32|f64] [ns] [nf]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "joss-nifty_b200"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import bench  # noqa: E402
import niftyb200 as nb  # noqa: E402

dt = sys.argv[1] if len(sys.argv) > 1 else "f32"
ns = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
nf = int(sys.argv[3]) if len(sys.argv) > 3 else 128
tdt = torch.float32 if dt == "f32" else torch.float64
s = 4 if dt == "f32" else 8
mk = nb.CorrelatedFieldMaker("", dtype=tdt)
mk.set_amplitude_total_offset(0.0, (1e-3, 1e-4))
mk.add_fluctuations((ns, ns), 1.0 / ns, (1e-1, 5e-3), (-1.0, 1e-2), (1.0, 0.5), (0.5, 5e-2), prefix="space",
                    non_parametric_kind="power")
mk.add_fluctuations((nf,), 1.0 / nf, (1.0, 1e-2), (-4.0, 1.0), (3.1, 1.0), (0.5, 0.1), prefix="freq",
                    non_parametric_kind="power")
cf = mk.finalize()
shape = cf.shape
rng = np.random.default_rng(43)
p = {k: rng.standard_normal(sh) for k, sh in cf.domain.items() if not k.endswith("xi")}
tab = cf.tables(p)
gen = torch.Generator(device="cuda").manual_seed(42)
w = torch.exp(0.3 * torch.randn(shape, dtype=tdt, device="cuda", generator=gen))
M = nb.CFMetric(cf, tab, w)
x0 = torch.randn(shape, dtype=tdt, device="cuda", generator=gen)
j = M(x0)
j.mul_(0.5).add_(x0)
cgm = sys.modules["niftyb200.cg"]
loop = cgm.CGLoop(M, j, x0)
loop.run(3)
N = int(np.prod(shape))
peak, _ = bench.peaks()
y = torch.empty_like(x0)


def tm(fn, reps=5):
    fn()
    return bench.time_region(torch, None, lambda: [fn() for _ in range(reps)]) / reps


t_it = bench.time_region(torch, None, lambda: loop.run(10)) / 10
t_f = tm(lambda: cf.apply_xi(x0, M.tab, out=y))
t_a = tm(lambda: cf.adjoint_xi(x0, M.tab, out=y))
d = 3
print(f"C5 {shape} {dt}: plans kinds {[pl.kind for pl in cf.plans]}  amp_mode {tab.amp_mode}")
print(f"  fwd {t_f:.3f} ms = {1000 / t_f:.1f} applies/s ({2 * d * N * s / t_f / 1e6 / peak:.3f} of nominal HBM roofline 2dNs)")
print(f"  adj {t_a:.3f} ms = {1000 / t_a:.1f} applies/s ({2 * d * N * s / t_a / 1e6 / peak:.3f})")
print(f"  CG iteration {t_it:.3f} ms = {1000 / t_it:.1f} it/s ({(4 * d + 13) * N * s / t_it / 1e6 / peak:.3f} of nominal (4d+13)Ns)")
