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


def run(rank, world, port, backend, emu_path, shape, result, mode="full", dtype="f64"):
    """mode "full": forward, adjoint, metric, 5 back-to-back transforms and a CG solve against the oracle;
    mode "large": the same without the CG solve (grids where the oracle CG would take minutes);
    mode "product": as "large" for the (axes 0, 1) x (axis 2) product of two spaces on the joint plan."""
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
        from oracle import cg as ocg, metric as om, transforms as otr
        cgm = sys.modules["niftyb200.cg"]
        dt = np.float64 if dtype == "f64" else np.float32
        tdt = torch.float64 if dtype == "f64" else torch.float32
        if rank == 0:
            otr.set_nthreads(max(1, (os.cpu_count() or 1) // 2))
        dist_ = tuple(1.0 / s for s in shape)      # isotropic harmonic pixels
        if mode == "product":
            cf, co = pc.make_product_pair([(shape[:2], dist_[:2]), (shape[2:], 0.7)], tdt, kw=pc.GENTLE_KW)
        else:
            cf, co = pc.make_pair(shape, dist_, tdt, kw=pc.GENTLE_KW)
        scf = nb.SlabCorrelatedField(cf)
        assert scf.product == (mode == "product")
        ht = scf.ht
        p = cf.init(5)
        p["xi"] = p["xi"].astype(dt)
        amps_o, azm_o = co.normalized_amplitudes(p), co.azm(p)
        tab = scf.streaming_tables(cf.tables_from_amplitudes(amps_o, float(azm_o)))
        if mode == "product":      # the slab tables straight from the latent parameters (no full outer product)
            t2 = scf.tables(p)
            assert float(torch.linalg.vector_norm((t2.amp - tab.amp).double())
                         / torch.linalg.vector_norm(tab.amp.double())) <= (1e-12 if dtype == "f64" else 1e-6)

        def f64(a):
            return np.asarray(a, dtype=np.float64)
        # forward apply: X -> Y
        y = scf.apply_xi(ht.scatter_x(p["xi"]), tab)
        got = ht.gather_y(y)
        err_f = pc.rel(got, co.apply(amps_o, azm_o, f64(p["xi"]))) if rank == 0 else 0.0
        # adjoint: Y -> X
        rng = np.random.default_rng(9)
        yb = rng.normal(size=shape).astype(dt)
        xb = scf.adjoint_xi(ht.scatter_y(yb), tab)
        got = ht.gather_x(xb)
        err_a = pc.rel(got, co.adjoint(amps_o, azm_o, f64(yb))[0]) if rank == 0 else 0.0
        # metric
        w = np.exp(0.3 * rng.normal(size=shape)).astype(dt)
        M = nb.CFMetric(scf, tab, ht.scatter_y(w))
        Mo = om.cf_metric_xi(co, amps_o, azm_o, f64(w))
        Mo.parts = (co, amps_o, azm_o, f64(w))
        x0 = rng.normal(size=shape).astype(dt)
        q, curv = M.apply_with_curvature(ht.scatter_x(x0))
        got = ht.gather_x(q)
        err_m = err_c = 0.0
        if rank == 0:
            ref = Mo(f64(x0))
            err_m = pc.rel(got, ref)
            err_c = abs(float(curv.cpu()) - np.vdot(f64(x0), ref)) / abs(np.vdot(f64(x0), ref))
        # double-buffer rule of the peer exchange: five back-to-back transforms, no host synchronisation
        # in between, compared after the last
        t = ht.scatter_x(x0)
        ys = [torch.empty(scf.position_shape, dtype=tdt, device=t.device) for _ in range(2)]
        xs = [torch.empty(scf.latent_shape, dtype=tdt, device=t.device) for _ in range(2)]
        y1 = scf.apply_xi(t, tab, out=ys[0], offset=0.0)
        x2 = scf.adjoint_xi(y1, tab, out=xs[0])
        y3 = scf.apply_xi(x2, tab, out=ys[1], offset=0.0)
        x4 = scf.adjoint_xi(y3, tab, out=xs[1])
        y5 = scf.apply_xi(x4, tab, out=ys[0], offset=0.0)
        got = ht.gather_y(y5)
        err_s = 0.0
        if rank == 0:
            r = f64(x0)
            for k in range(5):
                r = co.apply(amps_o, azm_o, r, offset=0.0) if k % 2 == 0 else co.adjoint(amps_o, azm_o, r)[0]
            err_s = pc.rel(got, r)
        out = dict(err_f=err_f, err_a=err_a, err_m=err_m, err_c=err_c, err_s=err_s, dtype=dtype, world=world,
                   p2p=ht.peer is not None)
        if mode == "full":
            j = om.linear_sample_rhs(co, amps_o, azm_o, f64(w), rng.normal(size=shape), f64(x0)).astype(dt)
            kw = dict(absdelta=1e-5 * j.size, maxiter=60)
            r = cgm._cg(M, ht.scatter_x(j), ht.scatter_x(x0), **kw)
            # device-resident static_cg on the slab-decomposed vectors (sums all-reduced on the stream) against
            # its host-driven twin: identical counts on every rank
            sd = cgm._static_cg(M, ht.scatter_x(j), ht.scatter_x(x0), **kw)
            sh = cgm._static_cg(M, ht.scatter_x(j), ht.scatter_x(x0), device_resident=False, **kw)
            assert tuple(sd[1:]) == tuple(sh[1:]), (rank, tuple(sd[1:]), tuple(sh[1:]))
            out.update(static=(sd.nit, sd.info, sd.nfev), err_static=float(torch.linalg.vector_norm(sd.x - sh.x)
                                                                         / torch.linalg.vector_norm(sh.x)))
            got = ht.gather_x(r.x)
            if rank == 0:
                ro = ocg.cg_re(Mo, f64(j), f64(x0), **kw)
                dE, g, nE, nG = pc.oracle_histories(Mo, f64(j), f64(x0), kw["maxiter"], dt)
                out.update(window=sorted(pc.acceptable_stops(dE, nE, kw["absdelta"], 6, kw["maxiter"])),
                           nit=(r.nit, ro.nit), info=(r.info, ro.info), nfev=(r.nfev, ro.nfev), err_x=pc.rel(got, ro.x))
        if rank == 0:
            result.update(out)
    finally:
        dist.destroy_process_group()


def launch(world, backend, emu_path, shape, port=29533, mode="full", dtype="f64"):
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    result = mgr.dict()
    mp.spawn(run, args=(world, port, backend, emu_path, shape, result, mode, dtype), nprocs=world, join=True)
    return dict(result)


def run_sampled(rank, world, port, emu_path, shape, result):
    """Samples sharded over the ranks (the reference's MPI sample parallelism, kl_energies.py:136-141) with one
    all-reduce of the N-vector per apply, against the same SampledCFMetric holding every sample on one rank."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    from niftyb200 import _lib
    _lib.use_library_for_testing(emu_path, "cpu")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import niftyb200 as nb
        import parity_cases as pc
        from niftyb200.distributed import Comm
        cgm = sys.modules["niftyb200.cg"]
        cf, co = pc.make_pair(shape, tuple(1.0 / s for s in shape), torch.float64, kw=pc.GENTLE_KW)
        rng = np.random.default_rng(3)
        nsamp = 5
        tabs, ws = [], []
        for s in range(nsamp):
            tabs.append(cf.tables(cf.init(50 + s)))
            ws.append(pc.dev(np.exp(0.3 * rng.normal(size=shape))))
        t = pc.dev(rng.normal(size=shape))
        j = pc.dev(rng.normal(size=shape))
        mine = list(range(rank, nsamp, world))
        M = nb.SampledCFMetric(cf, [tabs[s] for s in mine], [ws[s] for s in mine], comm=Comm(), n_total=nsamp)
        Mall = nb.SampledCFMetric(cf, tabs, ws)
        q, curv = M.apply_with_curvature(t)
        qa, curva = Mall.apply_with_curvature(t)
        r = cgm._cg(M, j, None, resnorm=1e-8 * float(torch.linalg.vector_norm(j)), maxiter=100)
        ra = cgm._cg(Mall, j, None, resnorm=1e-8 * float(torch.linalg.vector_norm(j)), maxiter=100)
        if rank == 0:
            result.update(err=pc.rel(pc.host(q), pc.host(qa)), errc=abs(float(curv) - float(curva)) / abs(float(curva)),
                          nit=(r.nit, ra.nit), errx=pc.rel(pc.host(r.x), pc.host(ra.x)))
    finally:
        dist.destroy_process_group()


def launch_sampled(world, emu_path, shape, port=29591):
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    result = mgr.dict()
    mp.spawn(run_sampled, args=(world, port, emu_path, shape, result), nprocs=world, join=True)
    return dict(result)
