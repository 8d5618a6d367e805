"""Generic path on non-power-of-two grids: chirp-z (Bluestein) axes vs the direct O(n^2) sum (NB2_BLUESTEIN=0)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "joss-nifty_b200"))
import torch
from niftyb200.transforms import Plan

def tk(fn, reps=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

for shape in ((1000, 1000), (3000, 3000), (384, 384, 384), (100, 100, 100), (1 << 20,), (1000003,), (1 << 24,)):
    for dt in (torch.float32, torch.float64):
        a = torch.randn(shape, dtype=dt, device="cuda")
        pl = Plan(shape, None, dt)
        t = tk(lambda: pl.hartley(a))
        print(f"NB2_BLUESTEIN={os.environ.get('NB2_BLUESTEIN','1')} {shape} {str(dt)[6:]}: {t:9.3f} ms", flush=True)
        del a, pl
