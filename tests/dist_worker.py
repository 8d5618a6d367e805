"""Worker for the world_size>1 tests: slab-decomposed correlated field + metric CG against
the single-process oracle.  Runs on CPU (gloo + kernel emulator) or on GPUs (nccl + CUDA)."""

import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "joss-nifty_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def run(rank, world, port, backend, emu_path, shape, result):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    from niftyb200 import _lib
    if backend == "gloo":
        _lib.use_library_for_testing(emu_path, "cpu")
    else:
        torch.cuda.set_device(rank)
    dist.init_process_group(backend, rank=rank, world_size=world)
    try:
        import niftyb200 as nb
        import parity_cases as pc
        from oracle import cg as ocg
        cgm = sys.modules["niftyb200.cg"]
        dist_ = tuple(1.0 / s for s in shape)      # isotropic harmonic pixels
        cf, co = pc.make_pair(shape, dist_, torch.float64, kw=pc.GENTLE_KW)
        scf = nb.SlabCorrelatedField(cf)
        ht = scf.ht
        p = cf.init(5)
        tab = scf.tables(p)
        amps_o, azm_o = co.normalized_amplitudes(p), co.azm(p)
        # forward apply: X -> Y
        y = scf.apply_xi(ht.scatter_x(p["xi"]), tab)
        err_f = pc.rel(ht.gather_y(y), co(p))
        # adjoint: Y -> X
        rng = np.random.default_rng(9)
        yb = rng.normal(size=shape)
        xb = scf.adjoint_xi(ht.scatter_y(yb), tab)
        err_a = pc.rel(ht.gather_x(xb), co.adjoint(amps_o, azm_o, yb)[0])
        # metric + CG
        from oracle import metric as om
        w = np.exp(0.3 * rng.normal(size=shape))
        M = nb.CFMetric(scf, tab, ht.scatter_y(w))
        Mo = om.cf_metric_xi(co, amps_o, azm_o, w)
        x0 = rng.normal(size=shape)
        j = om.linear_sample_rhs(co, amps_o, azm_o, w, rng.normal(size=shape), x0)
        q, curv = M.apply_with_curvature(ht.scatter_x(x0))
        err_m = pc.rel(ht.gather_x(q), Mo(x0))
        err_c = abs(float(curv.cpu()) - np.vdot(x0, Mo(x0))) / abs(np.vdot(x0, Mo(x0)))
        kw = dict(absdelta=1e-5 * j.size, maxiter=60)
        r = cgm._cg(M, ht.scatter_x(j), ht.scatter_x(x0), **kw)
        ro = ocg.cg_re(Mo, j, x0, **kw)
        err_x = pc.rel(ht.gather_x(r.x), ro.x)
        window = [ro.nit] + [ocg.cg_re(Mo, pc.perturb(j, sd), x0, **kw).nit for sd in (1, 2, 3, 4)]
        if rank == 0:
            result.update(window=window, err_f=err_f, err_a=err_a, err_m=err_m, err_c=err_c, nit=(r.nit, ro.nit), info=(r.info, ro.info),
                          nfev=(r.nfev, ro.nfev), err_x=err_x)
    finally:
        dist.destroy_process_group()


def launch(world, backend, emu_path, shape, port=29533):
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    result = mgr.dict()
    mp.spawn(run, args=(world, port, backend, emu_path, shape, result), nprocs=world, join=True)
    return dict(result)
