import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "joss-nifty_b200"))
import torch
from niftyb200.transforms import Plan, PointOp
shape = tuple(int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "2048,2048,128").split(","))
a = torch.randn(shape, dtype=torch.float32, device="cuda"); b = torch.empty_like(a)
w = torch.rand(shape, dtype=torch.float32, device="cuda") + 0.5
pl = Plan(shape, None, torch.float32)
pl.run_pass(2, a, b, epi=PointOp(scale=0.5, mul=w), product_last=True)
torch.cuda.synchronize()
torch.cuda.profiler.start()
pl.run_pass(2, a, b, epi=PointOp(scale=0.5, mul=w), product_last=True)
pl.run_pass(2, a, b)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
