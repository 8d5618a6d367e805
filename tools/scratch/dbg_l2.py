import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "joss-nifty_b200"))
import torch, niftyb200 as nb
n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
x = torch.randn(n, n, n, device="cuda"); y = torch.empty_like(x)
plan = nb.Plan((n, n, n), None, torch.float32)
os.environ["NB2_L2_STREAMS"] = "2"
for v in sys.argv[2].split(","):
    os.environ["NB2_L2_PLANES"] = v
    print("planes", v, flush=True)
    plan.hartley(x, out=y); torch.cuda.synchronize()
    plan.hartley(x, out=y); torch.cuda.synchronize()
    print("ok", flush=True)
