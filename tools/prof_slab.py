"""Timing breakdown of the slab-decomposed metric apply (run under torchrun).
Usage: torchrun ... tools/prof_slab.py [n] [dtype]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "joss-nifty_b200"))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import bench  # noqa: E402
import niftyb200 as nb  # noqa: E402
from niftyb200.transforms import PointOp  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
tdt = torch.float32 if (len(sys.argv) < 3 or sys.argv[2] == "f32") else torch.float64
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cf, tab, M, j, x0 = bench.setup_problem(nb, torch, [((n, n, n), bench.CF_KW)], bench.OFFSET, "gaussian", tdt, seed=42 + rank,
                                        slab=True)
ht = cf.ht
cgm = sys.modules["niftyb200.cg"]
loop = cgm.CGLoop(M, j, x0)
loop.run(3)
y = torch.empty(cf.position_shape, dtype=tdt, device="cuda")
xo = torch.empty_like(x0)


def tm(fn, reps=10):
    fn()
    return bench.time_region(torch, dist, lambda: [fn() for _ in range(reps)]) / reps


mc = ht.mc
res = {}
res["cg_iter"] = bench.time_region(torch, dist, lambda: loop.run(10)) / 10
res["metric_apply"] = tm(lambda: M.apply_with_curvature(x0))
res["x_to_y"] = tm(lambda: ht.x_to_y(x0, y))
res["y_to_x"] = tm(lambda: ht.y_to_x(y, xo))
if ht.peer is None:
    res["passA_x(axis1->send)"] = tm(lambda: [ht.plan_x.run_strided_remap(1, x0, ht.send[c], None, out_rowoff=ht.rowoff1,
                                              out_ostride=ht.ostride1, col0=c * mc, ncols=mc) for c in range(ht.chunks)])
    res["a2a_all_chunks"] = tm(lambda: [dist.all_to_all_single(ht.recv[c], ht.send[c]) for c in range(ht.chunks)])
    res["passB_y(recv->axis0)"] = tm(lambda: [ht.plan_y.run_strided_remap(0, ht.recv[c], y, in_rowoff=ht.rowoff0, col0=c * mc,
                                              ncols=ht.n1 * mc, run=mc, run_stride=ht.m) for c in range(ht.chunks)])
    res["passA_y(axis0->send)"] = tm(lambda: [ht.plan_y.run_strided_remap(0, y, ht.send[c], None, out_rowoff=ht.rowoff0, col0=c * mc,
                                              ncols=ht.n1 * mc, run=mc, run_stride=ht.m) for c in range(ht.chunks)])
    res["passB_x(recv->axis1)"] = tm(lambda: [ht.plan_x.run_strided_remap(1, ht.recv[c], xo, in_rowoff=ht.rowoff1,
                                              in_ostride=ht.ostride1, col0=c * mc, ncols=mc) for c in range(ht.chunks)])
else:
    mine = ht.peer.mine(0)
    res["passA_x(axis1->peers)"] = tm(lambda: ht.plan_x.run_strided_scatter(1, x0, mine, None, out_rowoff=ht.p2p_x[0],
                                                                            out_outer_off=ht.p2p_x_outer))
    res["arrive"] = tm(lambda: ht.peer.arrive())
    res["passB_y(mine->axis0)"] = tm(lambda: ht.plan_y.run_pass(0, mine, y))
    res["passA_y(axis0->peers)"] = tm(lambda: ht.plan_y.run_strided_scatter(0, y, mine, None, out_rowoff=ht.p2p_y[0],
                                                                            out_col_off=ht.p2p_y_col))
    res["passB_x(mine->axis1)"] = tm(lambda: ht.plan_x.run_pass(1, mine, xo))
res["rows_y"] = tm(lambda: ht.plan_y.run_pass(2, y, y))
res["rows_x"] = tm(lambda: ht.plan_x.run_pass(2, xo, xo))
res["cg_update"] = tm(lambda: loop.vec.update(0, 0.0, loop.d, loop.q, j, loop.pos, loop.r))
res["axpby"] = tm(lambda: loop.vec.axpby(1.0, loop.r, 1.0, xo))
t = torch.zeros(4, dtype=torch.float64, device="cuda")
res["allreduce_4xf64"] = tm(lambda: dist.all_reduce(t))
if rank == 0:
    tag = f"world={world} p2p={int(ht.peer is not None)} logw={os.environ.get('NB2_STRIDED_LOGW','-')} chunks={ht.chunks} hiprio={os.environ.get('TORCH_NCCL_HIGH_PRIORITY', '0')}"
    print(tag + "  " + "  ".join(f"{k}={v:.3f}" for k, v in res.items()), file=sys.stderr, flush=True)
dist.destroy_process_group()
