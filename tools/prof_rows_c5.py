"""Row pass of the c5 grid (2048 x 2048 x 128: rows of 64 complex) -- plain, product combine, metric epilogues."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "joss-nifty_b200"))
import torch
from niftyb200.transforms import Plan, PointOp

def tk(fn, reps=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

shape = tuple(int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "2048,2048,128").split(","))
DT = torch.float64 if (len(sys.argv) > 2 and sys.argv[2] == "f64") else torch.float32
a = torch.randn(shape, dtype=DT, device="cuda"); b = torch.empty_like(a)
f = torch.rand(shape, dtype=DT, device="cuda") + 0.5
w = torch.rand(shape, dtype=DT, device="cuda") + 0.5
curv = torch.zeros((), dtype=torch.float64, device="cuda")
pl = Plan(shape, None, torch.float32)
byt = 2 * a.numel() * a.element_size()
for prod in (False, True):
    for name, epi in (("plain", None), ("+weight", PointOp(scale=0.5, mul=w)),
                      ("+field+add+dot", PointOp(2, f, 0.5, add=w, dot=w, dot_out=curv, amp_sym=True))):
        t = tk(lambda: pl.run_pass(2, a, b, epi=epi, product_last=prod))
        print(f"prod={int(prod)} rows{name:16s} {t:7.3f} ms {byt / t / 1e6:7.0f} GB/s (2Ns)", flush=True)
