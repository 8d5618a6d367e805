"""Cluster strided pass (2048- / 4096-point axes) against StridedPass (NB2_CLUSTER=0), per pass."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "joss-nifty_b200"))
import torch  # noqa: E402
from niftyb200.transforms import Plan, PointOp  # noqa: E402


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


for shape in ((2048, 2048, 128), (4096, 4096), (2048, 64, 1024)):
    a = torch.randn(shape, dtype=torch.float32, device="cuda")
    b = torch.empty_like(a)
    f = torch.rand(shape, dtype=torch.float32, device="cuda") + 0.5
    r = torch.randn(shape, dtype=torch.float32, device="cuda")
    pl = Plan(shape, None, torch.float32)
    ref = pl.hartley(a)
    byt = 2 * a.numel() * 4
    first = 3 - len(shape)
    for w in range(first, 2):
        t = tk(lambda: pl.run_pass(w, a, b))
        print(f"{shape} NB2_CLUSTER={os.environ.get('NB2_CLUSTER', '1')} strided{w}: {t:7.3f} ms {byt / t / 1e6:7.0f} GB/s", flush=True)
    t = tk(lambda: pl.run_pass(first, a, b, pro=PointOp(2, f, 0.7)))
    print(f"{shape} strided{first}+field: {t:7.3f} ms {byt / t / 1e6:7.0f} GB/s (2Ns)", flush=True)
    d = a.clone()
    t = tk(lambda: pl.run_pass(first, d, b, pro=PointOp(2, f, 0.7, pre_add=r, pre_scale=0.5, pre_out=d)))
    print(f"{shape} strided{first}+field+update: {t:7.3f} ms {byt / t / 1e6:7.0f} GB/s (2Ns)", flush=True)
    print("checksum", float(ref.double().abs().sum()), flush=True)
    del a, b, f, r, d, pl, ref
