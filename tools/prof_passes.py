"""Profiling driver: each kernel of the hot path on one grid (for ncu / timing).
Usage: python tools/prof_passes.py [f32|f64] [c2|c3|c4|<n> (cubic grid)] [reps]
With PROF_NCU=1 every case is launched exactly ONCE, in the order printed on stderr ("case: <name>"), between
cudaProfilerStart/Stop, so that the k-th profiled kernel launch is the k-th case (tools/ncu_dram_summary.py)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "joss-nifty_b200"))
import torch  # noqa: E402
import bench  # noqa: E402
import niftyb200 as nb  # noqa: E402
from niftyb200.transforms import PointOp  # noqa: E402

dt = sys.argv[1] if len(sys.argv) > 1 else "f32"
which = sys.argv[2] if len(sys.argv) > 2 else "c4"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
if which in bench.CONFIGS:
    subs, offset, lik, _ = bench.CONFIGS[which]
else:
    subs, offset, lik = [((int(which),) * 3, bench.CF_KW)], bench.OFFSET, "gaussian"
shape = tuple(n for sh, _ in subs for n in sh)
tdt = torch.float32 if dt == "f32" else torch.float64
cf, tab, M, j, x0 = bench.setup_problem(nb, torch, subs, offset, lik, tdt)
plan = cf.plans[0] if getattr(cf, "joint", None) is None else cf.joint
PL = plan is not cf.plans[0]    # product of two spaces on the joint plan: the row pass combines differently
first = 3 - len(shape)          # first strided pass of the transform
a, b = x0, torch.empty_like(x0)
cgm = sys.modules["niftyb200.cg"]
loop = cgm.CGLoop(M, j, x0)
pro = PointOp(tab.amp_mode, tab.amp, tab.azm)
epi_f = PointOp(scale=cf.dvol, add_scalar=0.0, mul=M.w)
curv = torch.zeros((), dtype=torch.float64, device="cuda")
epi_a = PointOp(tab.amp_mode, tab.amp, tab.azm * cf.dvol, add=x0, dot=x0, dot_out=curv)
only = os.environ.get("PROF_ONLY")
SYM = M.tab.sym and os.environ.get("PROF_SYM", "1") != "0"   # NB2_AMP_SYMMETRIC reads of the streamed field
fieldpro = dict(amp_mode=M.tab.amp_mode, amp=M.tab.amp, scale=M.tab.azm, amp_sym=SYM)
cases = [] if first else [("strided0", lambda: plan.run_pass(0, a, b))]
cases += [("strided1", lambda: plan.run_pass(1, a, b)),
         ("rows", lambda: plan.run_pass(2, a, b, product_last=PL)), (f"strided{first}+amp", lambda: plan.run_pass(first, a, b, pro=pro)),
         ("rows+mul", lambda: plan.run_pass(2, a, b, epi=epi_f, product_last=PL)),
         ("rows+amp+add+dot", lambda: plan.run_pass(2, a, b, epi=epi_a, product_last=PL)),
         (f"strided{first}+field", lambda: plan.run_pass(first, a, b, pro=PointOp(**fieldpro))),
         (f"strided{first}+field+update", lambda: plan.run_pass(first, a, b, pro=PointOp(**fieldpro, pre_add=j, pre_scale=0.5,
                                                                                       pre_out=a))),
         ("rows+field+add+dot", lambda: plan.run_pass(2, a, b, epi=PointOp(M.tab.amp_mode, M.tab.amp, M.tab.azm * cf.dvol,
                                                                             add=x0, dot=x0, dot_out=curv, amp_sym=SYM),
                                                        product_last=PL)),
         ("cg_update", lambda: loop.vec.update(0, 0.0, loop.d, loop.q, j, loop.pos, loop.r)),
         ("axpby", lambda: loop.vec.axpby(1.0, loop.r, 1.0, b))]
if os.environ.get("PROF_NCU"):
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
for name, fn in cases:
    if only and not any(name == o for o in only.split(',')):
        continue
    print("case:", name, file=sys.stderr, flush=True)
    fn()
    if reps == 0:
        torch.cuda.synchronize()
        continue
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    N = x0.numel()
    print(f"{name:20s} {ms:8.3f} ms   {2 * N * x0.element_size() / ms / 1e6:8.1f} GB/s (2Ns)", flush=True)
if os.environ.get("PROF_NCU"):
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
