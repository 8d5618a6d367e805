"""Timing of the L2-blocked tail: plain transform, metric apply and CG iteration for several
group sizes (NB2_L2_PLANES is read per call).  Usage: python tools/prof_l2.py [f32|f64] [n] [reps] [planes,...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "joss-nifty_b200"))
import torch  # noqa: E402
import bench  # noqa: E402
import niftyb200 as nb  # noqa: E402

dt = sys.argv[1] if len(sys.argv) > 1 else "f32"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
variants = sys.argv[4].split(",") if len(sys.argv) > 4 else ["0", "4", "8", "16", "32"]
shape = (n, n, n)
tdt = torch.float32 if dt == "f32" else torch.float64
cf, tab, M, j, x0 = bench.setup_problem(nb, torch, shape, tdt)
plan = cf.plans[0]
a, b = x0.clone(), torch.empty_like(x0)
cgm = sys.modules["niftyb200.cg"]
loop = cgm.CGLoop(M, j, x0)


def timed(fn):
    fn(); fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for v in variants:
    if "x" in v:     # planes x streams
        v, ns = v.split("x")
        os.environ["NB2_L2_STREAMS"] = ns
    if v == "auto":
        os.environ.pop("NB2_L2_PLANES", None)
    else:
        os.environ["NB2_L2_PLANES"] = v
    th = timed(lambda: plan.hartley(a, out=b))
    tm = timed(lambda: M(a, out=b))
    tc = timed(lambda: loop.run(1))
    print(f"planes={v:5s} streams={os.environ.get('NB2_L2_STREAMS', 'dflt')} hartley {th:7.3f} ms  metric apply {tm:7.3f} ms  CG iteration {tc:7.3f} ms", flush=True)
