import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "joss-nifty_b200"))
import torch
from niftyb200.transforms import Plan
shape = (2048, 2048, 128)
a = torch.randn(shape, dtype=torch.float32, device="cuda")
b = torch.empty_like(a)
pl = Plan(shape, None, torch.float32)
pl.run_pass(0, a, b)
torch.cuda.synchronize()
torch.cuda.profiler.start()
pl.run_pass(0, a, b)
pl.run_pass(1, a, b)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
