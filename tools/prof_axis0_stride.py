"""Axis-0 strided pass (TMA kernel, 1024 rows) as a function of the row stride: same rows, same tile, same
bytes per launch -- only the distance between consecutive rows of a tile changes (N1 * N2 * 4 bytes)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "joss-nifty_b200"))
import torch  # noqa: E402
from niftyb200.transforms import Plan  # noqa: E402


def tk(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for n1, n2 in ((16, 1024), (64, 1024), (256, 1024), (1024, 256), (1024, 1024), (2048, 512)):
    shape = (1024, n1, n2)
    a = torch.randn(shape, dtype=torch.float32, device="cuda")
    b = torch.empty_like(a)
    pl = Plan(shape, None, torch.float32)
    t0, t1 = tk(lambda: pl.run_pass(0, a, b)), tk(lambda: pl.run_pass(1, a, b))
    byt = 2 * a.numel() * 4
    print(f"{shape}: row stride {n1 * n2 * 4 / 1024:8.0f} KB  axis0 {t0:7.3f} ms {byt / t0 / 1e6:7.0f} GB/s   "
          f"axis1 {t1:7.3f} ms {byt / t1 / 1e6:7.0f} GB/s", flush=True)
    del a, b, pl
