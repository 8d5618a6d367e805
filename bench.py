#!/usr/bin/env python3
"""bench.py -- metric-CG iterations/s (each = 1 CF forward + 1 CF adjoint apply + fused CG
vector updates) on a synthetic correlated field, BASELINE.json's metric and config.

  python bench.py --gpus N --steps K --warmup W [--config c2|c3|c4|c5] [--dtype f32|f64] [--shape 1024,1024,1024]
  python bench.py --impl reference ...     # the reference's CPU algorithm on the host cores

--config selects a BASELINE.json configuration (default c4 = the 1024^3 grid the metric is quoted on):
  c2  4096 x 4096, Poissonian metric (w = exp(s)), demos/0_intro.py hyper-parameters
  c3  512^3, tomography hyper-parameters, Gaussian metric with a dense inverse variance
  c4  1024^3, same model (slab-decomposed when --gpus > 1)
  c5  2048 x 2048 space x 128 frequency product space (two amplitude spectra), Gaussian metric

One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for the definitions.
"""

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "joss-nifty_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

if "reference" in sys.argv:
    # the CPU arm uses every host core, also under torchrun (which exports OMP_NUM_THREADS=1);
    # must happen before NumPy loads its BLAS
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_v] = str(os.cpu_count() or 1)

import numpy as np  # noqa: E402

FUSED = "StridedPass(first axis)+amplitude field+direction update (fused axpby)"
G_STRIDED3 = "strided pass (StridedTma / StridedPass): first axis + amplitude field, axis 1, axis 0, axis 1"
G_STRIDED2 = "strided pass (StridedTma / StridedPass): axis 1 + amplitude field, axis 1"
G_ROWS = "row pass (RowPass): + weight, + amplitude field + add + curvature dot"
G_APPLY = "transform passes of the CF forward + adjoint apply (strided passes, mirror combine, row passes)"
NVLINK_GBS = 770.0   # GB/s per direction per GPU: measured peer copy (B200_PROFILING.md; nominal 900)
METRIC = "metric_cg_iters_per_s"
UNIT = "CG iterations/s"
# demos/1_tomography.py:21-34 hyper-parameters
CF_KW = dict(fluctuations=(1.0, 0.5), loglogavgslope=(-5.0, 0.2), flexibility=(1.0, 0.2), asperity=(0.5, 5e-2))
OFFSET = (2.0, (1e-1, 3e-2))
# demos/0_intro.py:24-37 (configs C1/C2) and the frequency axis of C5 (test_correlated_fields.py:159-165)
CF_KW_2D = dict(fluctuations=(1e-1, 5e-3), loglogavgslope=(-1.0, 1e-2), flexibility=(1.0, 0.5), asperity=(0.5, 5e-2))
OFFSET_2D = (0.0, (1e-3, 1e-4))
CF_KW_FREQ = dict(fluctuations=(1.0, 1e-2), loglogavgslope=(-4.0, 1.0), flexibility=(3.1, 1.0), asperity=(0.5, 0.1))

# name -> (sub-spaces [(shape, hyper-parameters)], offset, likelihood, description)
CONFIGS = {
    "c2": ([((4096, 4096), CF_KW_2D)], OFFSET_2D, "poisson",
           "2D correlated field 4096x4096, Poissonian metric-CG (configs[1])"),
    "c3": ([((512, 512, 512), CF_KW)], OFFSET, "gaussian", "3D correlated field 512x512x512 metric-CG (configs[2])"),
    "c4": ([((1024, 1024, 1024), CF_KW)], OFFSET, "gaussian", "3D correlated field 1024x1024x1024 metric-CG (configs[3] grid)"),
    "c5": ([((2048, 2048), CF_KW_2D), ((128,), CF_KW_FREQ)], OFFSET_2D, "gaussian",
           "product-space correlated field 2048x2048 space x 128 frequency metric-CG (configs[4] grid)"),
}


def resolve_config(args):
    """(sub-spaces, offset, likelihood, description, full shape) of the run."""
    subs, offset, lik, desc = CONFIGS[args.config]
    if args.shape:
        shape = tuple(int(v) for v in args.shape.split(","))
        subs = [(shape, subs[0][1])]
        desc = f"{len(shape)}D correlated field {'x'.join(map(str, shape))} metric-CG (--shape, {args.config} model)"
    shape = tuple(n for sh, _ in subs for n in sh)
    return subs, offset, lik, desc, shape


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--config", default="c4", choices=sorted(CONFIGS))
    ap.add_argument("--shape", default="", help="override the grid of a single-space config")
    ap.add_argument("--cpu-shape", default="", help="bounded sample for the cpu_baseline leg (default: per config)")
    ap.add_argument("--no-verify", action="store_true", help="skip the pre-timing parity check of multi-GPU runs")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def dram_traffic(config, dtype, group):
    """DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of the kernel group `group`,
    averaged over the launches of the group inside one CG iteration, from the committed summary of an
    `ncu --set full` capture of tools/prof_passes.py (profiles/dram_bytes.json, written by
    tools/ncu_dram_summary.py; per kernel case).  (None, reason) when that config was not captured."""
    path = os.path.join(ROOT, "profiles", "dram_bytes.json")
    try:
        with open(path) as f:
            db = json.load(f)
        cases = db[f"{config}/{dtype}"]["cases"]
    except Exception:
        return None, "no ncu capture committed for this config (profiles/dram_bytes.json)"
    want = {G_STRIDED3: ["strided0+field+update", "strided1", "strided0", "strided1"],
            G_STRIDED2: ["strided1+field+update", "strided1"],
            G_ROWS: ["rows+mul", "rows+field+add+dot"], "CgUpdate (+3 dots)": ["cg_update"], "Axpby": ["axpby"]}.get(group)
    if not want or any(w not in cases for w in want):
        return None, "kernel group not in profiles/dram_bytes.json"
    return sum(cases[w]["dram_bytes"] for w in want) / len(want), db[f"{config}/{dtype}"].get("source", path)


# ---------------------------------------------------------------- clocks during the timed region
class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU during the timed region.

    NVML in a thread (10 ms period; the same counters nvidia-smi prints, without its start-up
    latency, which is longer than a short multi-GPU timed region); `nvidia-smi -lms` as fallback."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index, self.stop, self.th = [], None, index, False, None

    def _nvml_handle(self):
        import pynvml
        import torch
        pynvml.nvmlInit()
        try:
            uuid = "GPU-" + str(torch.cuda.get_device_properties(self.index).uuid)
            return pynvml, pynvml.nvmlDeviceGetHandleByUUID(uuid)
        except Exception:
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.index)

    def _poll(self, nv, h):
        bits = (("hw_slowdown", nv.nvmlClocksThrottleReasonHwSlowdown),
                ("hw_thermal_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown),
                ("sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwThermalSlowdown),
                ("sw_power_cap", nv.nvmlClocksThrottleReasonSwPowerCap))
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        while not self.stop:
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                self.rows.append([str(sm), str(mx)] + ["Active" if mask & b else "Not Active" for _, b in bits])
            except Exception:
                pass
            time.sleep(0.01)

    def __enter__(self):
        try:
            nv, h = self._nvml_handle()
            self.th = threading.Thread(target=self._poll, args=(nv, h), daemon=True)
            self.th.start()
            return self
        except Exception:
            self.th = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        self.stop = True
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        elif self.th is not None:
            self.th.join(timeout=1)

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------- CPU arm: the reference's algorithm on host cores
def tune_blas_threads(cores):
    """The CPU arm should be as fast as this host allows: the FFTs use every core (scipy workers); for the level-1
    BLAS behind np.vdot the fastest of {all cores, 1 thread} is taken -- on some containers a multi-threaded ddot
    is 100x slower than a single-threaded one (threads spinning against the FFT pool).  Returns the count used."""
    try:
        from threadpoolctl import threadpool_limits
    except Exception:
        return None
    a = np.ones(1 << 24)
    best, best_t = cores, None
    for n in (cores, 1):
        threadpool_limits(limits=n)
        np.vdot(a, a)
        t0 = time.perf_counter()
        np.vdot(a, a)
        t = time.perf_counter() - t0
        if best_t is None or t < best_t:
            best, best_t = n, t
    threadpool_limits(limits=best)
    return best


def cpu_cg_iters_per_s(subs, offset, lik, dtype, iters=1):
    """Oracle (NumPy/SciPy restatement of the reference, oracle/) CG iterations on the host cores: the
    reference's fallback path, scipy pocketfft with workers = all cores.  `iters` CG iterations from x0 = 0
    (no set-up apply: re/conjugate_gradient.py:108-112), each = one metric apply (CF forward + adjoint) +
    the vector updates and dots.  Returns (iterations/s, cores, seconds)."""
    from oracle import correlated_field as ocf, metric as om, cg as ocg, transforms as otr
    cores = os.cpu_count() or 1
    otr.set_nthreads(cores)
    tune_blas_threads(cores)
    dt = np.float32 if dtype == "f32" else np.float64
    mk = ocf.CorrelatedFieldMaker("")
    mk.set_amplitude_total_offset(*offset)
    for i, (shape, kw) in enumerate(subs):
        mk.add_fluctuations(shape, 1.0 / shape[0], **kw, prefix="" if len(subs) == 1 else f"ax{i}",
                            non_parametric_kind="power")
    cf = mk.finalize()
    rng = np.random.default_rng(43)
    p = {k: rng.standard_normal(s) for k, s in cf.domain.items() if k != "xi"}
    amps, azm = cf.normalized_amplitudes(p), cf.azm(p)
    rng = np.random.default_rng(42)
    j = rng.standard_normal(cf.shape, dtype=dt)
    # the values of the weight do not change the cost of an iteration (Poissonian: w = exp(s) > 0 as well)
    w = np.exp(dt(0.3) * rng.standard_normal(cf.shape, dtype=dt))
    mat = om.cf_metric_xi(cf, [a.astype(dt) for a in amps], dt(azm), w)
    t0 = time.perf_counter()
    res = ocg.cg_re(mat, j, None, absdelta=0.0, miniter=iters, maxiter=iters, _raise_nonposdef=False)
    t = time.perf_counter() - t0
    return res.nit / t, cores, t


def host_memory_gb():
    try:
        import psutil
        return psutil.virtual_memory().available / 2 ** 30
    except Exception:
        return 0.0


def cpu_sample_subs(subs, want_shape=""):
    """Bounded sample of the workload for the cpu_baseline leg of the default run (10-30 s of CPU work):
    the same model on a grid of at most 2^27 points, halving every axis of the first sub-space until it fits."""
    if want_shape:
        shape = tuple(int(v) for v in want_shape.split(","))
        return [(shape, subs[0][1])] + list(subs[1:])
    out = [(tuple(sh), kw) for sh, kw in subs]
    while np.prod([n for sh, _ in out for n in sh]) > 2 ** 27:
        sh, kw = out[0]
        out[0] = (tuple(max(2, n // 2) for n in sh), kw)
    return out


def scale_to(shape_from, shape_to):
    """N log N work ratio used to express a bounded CPU sample in the headline workload's unit."""
    nf, nt = float(np.prod(shape_from)), float(np.prod(shape_to))
    return (nf * np.log2(nf)) / (nt * np.log2(nt))


# ---------------------------------------------------------------- GPU arm
def make_field(nb, subs, offset, tdt):
    mk = nb.CorrelatedFieldMaker("", dtype=tdt)
    mk.set_amplitude_total_offset(*offset)
    for i, (shape, kw) in enumerate(subs):
        mk.add_fluctuations(shape, 1.0 / shape[0], **kw, prefix="" if len(subs) == 1 else f"ax{i}",
                            non_parametric_kind="power")
    return mk.finalize()


def setup_problem(nb, torch, subs, offset, lik, tdt, seed=42, slab=False):
    """Synthetic problem: xi ~ N(0,1), the config's hyper-parameters, spectrum latents ~ N(0,1)
    (seed 43); likelihood weight: Gaussian = dense per-voxel inverse variance, Poissonian = exp(cf(position))
    (likelihood_impl.py:248-258).  slab=True: the grid slab-decomposed over the ranks (local parts of the
    weight in layout Y and of the vectors in layout X)."""
    cf0 = make_field(nb, subs, offset, tdt)
    cf = nb.SlabCorrelatedField(cf0) if slab else cf0
    rng = np.random.default_rng(43)
    p = {k: rng.standard_normal(s) for k, s in cf0.domain.items() if not k.endswith("xi")}
    tab = cf.tables(p)
    gen = torch.Generator(device="cuda").manual_seed(seed)
    pshape = cf.position_shape if slab else cf0.shape
    lshape = cf.latent_shape if slab else cf0.shape
    if lik == "poisson":
        pos = torch.randn(lshape, dtype=tdt, device="cuda", generator=gen)
        w = torch.exp(cf.apply_xi(pos, tab))
        del pos
    else:
        w = torch.exp(0.3 * torch.randn(pshape, dtype=tdt, device="cuda", generator=gen))
    M = nb.CFMetric(cf, tab, w)
    x0 = torch.randn(lshape, dtype=tdt, device="cuda", generator=gen)
    j = M(x0)
    j.mul_(0.5).add_(x0)
    return cf, tab, M, j, x0


def verify_slab(nb, torch, dist, rank, world, vsubs=None, voffset=OFFSET):
    """Pre-timing parity check of a multi-GPU run, OUTSIDE the timed region: a 256^3 (product spaces:
    256 x 256 x 64) fp64 metric apply through the same SlabCorrelatedField / PeerExchange code path the timed
    loop uses, gathered and compared on rank 0 with the CPU oracle (the checker: oracle.metric.cf_metric_xi).
    Returns the rel-L2 error."""
    from niftyb200.transforms import device as nb_device
    dev = nb_device()
    vsubs = [((256, 256, 256), CF_KW)] if vsubs is None else vsubs
    shape = tuple(n for sh, _ in vsubs for n in sh)
    cf0 = make_field(nb, vsubs, voffset, torch.float64)
    cf = nb.SlabCorrelatedField(cf0)
    rng = np.random.default_rng(43)
    p = {k: rng.standard_normal(s) for k, s in cf0.domain.items() if not k.endswith("xi")}
    rng = np.random.default_rng(7)
    w = np.exp(0.3 * rng.standard_normal(shape))
    t = rng.standard_normal(shape)
    err = torch.zeros(1, dtype=torch.float64, device=dev)
    if rank == 0:
        from oracle import correlated_field as ocf, metric as om, transforms as otr
        otr.set_nthreads(max(1, (os.cpu_count() or 2) // 2))
        mk = ocf.CorrelatedFieldMaker("")
        mk.set_amplitude_total_offset(*voffset)
        for i, (sh, kw) in enumerate(vsubs):
            mk.add_fluctuations(sh, 1.0 / sh[0], **kw, prefix="" if len(vsubs) == 1 else f"ax{i}",
                                non_parametric_kind="power")
        co = mk.finalize()
        amps, azm = co.normalized_amplitudes(p), co.azm(p)
        ref = om.cf_metric_xi(co, amps, azm, w)(t)
        otr.set_nthreads(1)
        bc = [amps, float(azm)]
    else:
        ref, bc = None, [None, None]
    dist.broadcast_object_list(bc, src=0)       # the same amplitudes on both sides (and on every rank)
    tab = cf.streaming_tables(cf0.tables_from_amplitudes(bc[0], bc[1]))
    M = nb.CFMetric(cf, tab, cf.ht.scatter_y(w))
    q, _ = M.apply_with_curvature(cf.ht.scatter_x(t))
    q2, _ = M.apply_with_curvature(cf.ht.scatter_x(t))      # second buffer of the exchange
    got, got2 = cf.ht.gather_x(q), cf.ht.gather_x(q2)
    if rank == 0:
        nrm = np.linalg.norm(ref)
        err[0] = max(np.linalg.norm(got - ref), np.linalg.norm(got2 - ref)) / nrm
    dist.broadcast(err, src=0)
    cf.close()
    del M, tab, cf, cf0
    if dev.type == "cuda":
        torch.cuda.empty_cache()
    return float(err.item())


def time_region(torch, dist, fn):
    """barrier + sync, CUDA events on the launching stream, barrier + sync; max over ranks."""
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if dist is not None:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        dist.barrier()
    return ms


def run_b200(args):
    import torch
    import niftyb200 as nb
    from niftyb200 import _lib
    from niftyb200.transforms import PointOp
    cgm = sys.modules["niftyb200.cg"]
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist_mod.init_process_group("nccl", device_id=torch.device("cuda", local))
        dist = dist_mod
    lib = _lib.get()
    subs, offset, lik, desc, shape = resolve_config(args)
    tdt = torch.float32 if args.dtype == "f32" else torch.float64
    s = 4 if args.dtype == "f32" else 8
    N, d = int(np.prod(shape)), len(shape)
    peak, peak_src = peaks()

    parity = None
    if world > 1:
        product = len(subs) == 2 and len(subs[0][0]) == 2 and len(subs[1][0]) == 1
        if len(shape) != 3 or not (len(subs) == 1 or product):
            raise SystemExit("bench.py: --gpus > 1 shards one 3-D grid (c3, c4, c5); c2 runs on one GPU")
        if not args.no_verify:
            vs = [((256, 256), subs[0][1]), ((64,), subs[1][1])] if product else None
            err = verify_slab(nb, torch, dist, rank, world, vs, offset if product else OFFSET)
            parity = {"checked": True, "rel_l2": err, "ranks": world, "tolerance": 1e-12,
                      "what": ("256x256 space x 64 frequency" if product else "256^3") +
                              " fp64 metric apply (two consecutive, both exchange buffers) through the same slab / "
                              "peer-exchange code path, gathered, vs the CPU oracle, before the timed region"}
            if not err <= 1e-12:
                if rank == 0:
                    sys.stderr.write(f"bench.py: multi-GPU parity check FAILED: rel-L2 {err:.3e} > 1e-12\n")
                dist.destroy_process_group()
                sys.exit(3)
        cf, tab, M, j, x0 = setup_problem(nb, torch, subs, offset, lik, tdt, seed=42 + rank, slab=True)
    else:
        cf, tab, M, j, x0 = setup_problem(nb, torch, subs, offset, lik, tdt, seed=42)
    loop = cgm.CGLoop(M, j, x0)
    loop.run(args.warmup)
    n0, g0 = lib.nb2_launch_count(), loop.replayed_kernels
    with ClockSampler(local) as clk:
        ms = time_region(torch, dist, lambda: loop.run(args.steps))
    # kernels launched one by one + kernels inside replayed CUDA graphs (counted when the graph was captured)
    launches = int(lib.nb2_launch_count() - n0) + int(loop.replayed_kernels - g0)
    ms_step = ms / args.steps
    value = 1000.0 / ms_step                  # ONE global grid, slab-decomposed over the ranks when world > 1

    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
           # ONE grid of fixed size whatever N: the total work does not grow with the number of GPUs
           "scaling": "strong",
           "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
           "config": {"workload": desc + "; 1 step = 1 CG iteration = CF fwd + CF adj + fused updates",
                      "name": args.config, "likelihood": lik,
                      "loop": ("device-resident CG loop (scalars, alpha/beta and stopping tests on the device, nothing read "
                               "back inside the timed region" + ("; regular iteration replayed as a CUDA graph)"
                                                                 if getattr(loop.dev, "use_graph", False) else ")"))
                      if loop.dev is not None else "host-driven CG loop (4 scalars read back per iteration)",
                      "l2": "inputs larger than L2 (each vector %.2f GB per GPU)" % (N * s / 1e9 / world),
                      "multi_gpu": (f"slab-decomposed over {world} GPUs; exchange fused into the strided pass before it: "
                                    "every output row is stored straight into its owner's HBM over NVLink (CUDA IPC peer "
                                    "memory, nb2_strided_pass_scatter), one NCCL all-reduce of a scalar per transform as the "
                                    "arrival barrier, NCCL all-reduce of <= 3 fp64 scalars per inner-product group"
                                    if getattr(cf.ht, "peer", None) is not None else
                                    f"slab-decomposed over {world} GPUs; NCCL all_to_all_single per transform (column chunks), "
                                    "all-reduce of <= 3 fp64 scalars per inner-product group")
                      if world > 1 else "single GPU"},
           "gpu_launches": launches, "clocks": clk.summary()}
    if parity is not None:
        out["parity"] = parity
    B_cg = (4 * d + 13) * N * s                # SURVEY.md 8(d) nominal bytes per CG iteration
    out["extra"] = {"cg_nominal_bytes": B_cg, "cg_hbm_frac_of_measured": B_cg / world / (ms_step * 1e-3) / 1e9 / peak}
    if world > 1:
        # combined roofline: max(HBM time of the local share, NVLink time of two all-to-alls)
        nvl = NVLINK_GBS
        t_hbm = B_cg / world / (peak * 1e9)
        t_nvl = 2 * (world - 1) / world * N * s / world / (nvl * 1e9)
        out["extra"].update({"combined_roofline_ms": max(t_hbm, t_nvl) * 1e3, "t_hbm_ms": t_hbm * 1e3, "t_nvlink_ms": t_nvl * 1e3,
                             "frac_of_combined_roofline": max(t_hbm, t_nvl) * 1e3 / ms_step, "nvlink_gbs_assumed": nvl})

    # ---- forward / adjoint applies per second (every N; collective when slab-decomposed)
    reps = max(3, min(10, args.steps))
    a, b = x0, torch.empty_like(x0)
    yb = torch.empty(M.position_shape, dtype=tdt, device="cuda")
    tabs = M.tab          # amplitude gather resolved once per expansion point (streamed field)

    def tk(fn, d_=None):
        fn()
        return time_region(torch, d_, lambda: [fn() for _ in range(reps)]) / reps

    t_fwd = tk(lambda: cf.apply_xi(a, tabs, out=yb), dist)
    t_adj = tk(lambda: cf.adjoint_xi(yb, tabs, out=b), dist)
    roof_apply = 2 * d * N * s / world / (peak * 1e9) * 1e3      # ms, HBM term of one apply
    if world > 1:
        roof_apply = max(roof_apply, (world - 1) / world * N * s / world / (NVLINK_GBS * 1e9) * 1e3)
    out["extra"].update({"fwd_applies_per_s": 1000.0 / t_fwd, "adj_applies_per_s": 1000.0 / t_adj,
                         "fwd_adj_pairs_per_s": 1000.0 / (t_fwd + t_adj),
                         "fwd_frac_of_roofline": roof_apply / t_fwd, "adj_frac_of_roofline": roof_apply / t_adj})
    del yb

    # ---- end to end through the public API with HOST buffers (every rank moves its local part)
    if not args.no_e2e:
        lshape = tuple(j.shape)
        jh = torch.empty(lshape, dtype=tdt, pin_memory=True).copy_(j)
        xh = torch.empty(lshape, dtype=tdt, pin_memory=True).copy_(x0)
        oh = torch.empty(lshape, dtype=tdt, pin_memory=True)
        K = args.steps

        def solve():
            jd, xd = jh.to("cuda", non_blocking=True), xh.to("cuda", non_blocking=True)
            xs, info = nb.cg(M, jd, xd, absdelta=0.0, miniter=K, maxiter=K, _raise_nonposdef=False)
            oh.copy_(xs, non_blocking=True)
            torch.cuda.synchronize()
            return info

        solve()
        if dist is not None:
            dist.barrier()
        t0 = time.perf_counter()
        solve()
        if dist is not None:
            dist.barrier()
        te = time.perf_counter() - t0
        if dist is not None:
            tt = torch.tensor([te], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            te = float(tt.item())
        out["e2e"] = {"value": K / te, "unit": UNIT, "h2d_bytes_per_step": 2 * N * s / K,
                      "d2h_bytes_per_step": N * s / K + 24 * world,
                      "what": f"nb.cg(...) solve of {K} iterations, j and x0 from pinned host memory, "
                              "solution back to host, per-iteration scalars read by the host"
                              + (f"; every one of the {world} ranks moves its slab" if world > 1 else "")}
        del jh, xh, oh

    if rank == 0 and world > 1:
        ach = B_cg / world / (ms_step * 1e-3) / 1e9
        out["roofline"] = {"bound": "hbm", "kernel": "whole CG iteration, local share (9 kernels incl. the 2 passes whose "
                                                     "stores cross NVLink)",
                           "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                           "peak_source": peak_src, "algorithmic_bytes_per_launch": B_cg / world}
        emit(out)
    if rank == 0 and world == 1:
        # ---- per-kernel timing (CUDA events, same stream) and the roofline of the dominant kernel
        plan = cf.plans[0] if getattr(cf, "joint", None) is None else cf.joint
        pl_ = plan is not cf.plans[0]          # product of two spaces on one joint plan (NB2_HARTLEY_PRODUCT_LAST)
        t_fwd_g = tk(lambda: cf.apply_xi(a, tab, out=b)) if tab is not tabs else t_fwd
        t_adj_g = tk(lambda: cf.adjoint_xi(a, tab, out=b)) if tab is not tabs else t_adj
        names = {0: "StridedPass(axis0)", 1: "StridedPass(axis1)", 2: "RowPass(contiguous+Hartley combine)"}
        passes = {}
        if plan.is_fast and (cf.single or pl_):
            pro = PointOp(tab.amp_mode, tab.amp, tab.azm)
            for which in range(3):
                if len(shape) - 3 + which < 0:
                    continue
                passes[names[which]] = tk(lambda w=which: plan.run_pass(w, a, b, product_last=pl_))
            if len(shape) > 1:
                pros = PointOp(tabs.amp_mode, tabs.amp, tabs.azm, amp_sym=tabs.sym)
                passes["StridedPass(first axis)+amplitude field"] = tk(lambda: plan.run_pass(3 - len(shape), a, b, pro=pros))
                passes["StridedPass(first axis)+amplitude gather"] = tk(lambda: plan.run_pass(3 - len(shape), a, b, pro=pro))
                if getattr(M, "fused_direction_update", False):
                    # the first pass of the metric apply as the CG loop runs it: d <- r + beta d in place, x field
                    prod = PointOp(tabs.amp_mode, tabs.amp, tabs.azm, amp_sym=tabs.sym, pre_add=j, pre_scale=0.5, pre_out=a)
                    passes[FUSED] = tk(lambda: plan.run_pass(3 - len(shape), a, b, pro=prod))
            passes["RowPass+weight (metric fwd epilogue)"] = tk(lambda: plan.run_pass(
                2, a, b, epi=PointOp(scale=cf.dvol, mul=M.w), product_last=pl_))
            passes["RowPass+amplitude field+add+dot (metric adj epilogue)"] = tk(lambda: plan.run_pass(
                2, a, b, epi=PointOp(tabs.amp_mode, tabs.amp, tabs.azm * cf.dvol, add=x0, dot=x0, dot_out=M._curv,
                                     amp_sym=tabs.sym), product_last=pl_))
        t_upd = tk(lambda: loop.vec.update(0, 0.0, loop.d, loop.q, j, loop.pos, loop.r))
        t_axp = tk(lambda: loop.vec.axpby(1.0, loop.r, 1.0, b))
        # the kernels that run inside one CG iteration (the gather variant is reported, not used there)
        # ALGORITHMIC bytes per launch: every full-size operand once
        # (the curvature dot re-reads the array that is added, so add+dot count once)
        kbytes = {k: (2 + ("field" in k) + ("weight" in k) + ("add" in k) + 2 * ("direction update" in k)) * N * s
                  for k in passes}
        # The launches of ONE CG iteration grouped by kernel function; the dominant kernel is the one with
        # the largest share of the step.  achieved = algorithmic bytes per launch / average launch duration.
        fused = bool(passes.get(FUSED))
        f0 = FUSED if fused else "StridedPass(first axis)+amplitude field"
        groups = {}
        if len(shape) == 3 and all(passes.get(k) for k in (f0, names[0], names[1])):
            groups[G_STRIDED3] = [f0, names[1], names[0], names[1]]
        elif len(shape) == 2 and passes.get(f0) and passes.get(names[1]):
            groups[G_STRIDED2] = [f0, names[1]]
        rk = ["RowPass+weight (metric fwd epilogue)", "RowPass+amplitude field+add+dot (metric adj epilogue)"]
        if all(passes.get(k) for k in rk):
            groups[G_ROWS] = rk
        gsum = {g: (sum(passes[k] for k in ks), sum(kbytes[k] for k in ks), len(ks)) for g, ks in groups.items()}
        if not groups:
            # product spaces (one transform per sub-space, chained): the passes of both applies as one group
            nl = 2 * sum(2 if len(sh) > 1 else 1 for sh, _ in subs) + 2
            gsum[G_APPLY] = (t_fwd + t_adj, (4 * d + 3) * N * s, nl)
        gsum["CgUpdate (+3 dots)"] = (t_upd, 7 * N * s, 1)
        if not fused:       # otherwise the direction update rides on the first strided pass
            gsum["Axpby"] = (t_axp, 3 * N * s, 1)
        dom = max(gsum, key=lambda g: gsum[g][0])
        t_dom, b_dom, n_dom = gsum[dom]
        ach = b_dom / (t_dom * 1e-3) / 1e9
        traffic, traffic_src = dram_traffic(args.config if not args.shape else "", args.dtype, dom)
        out["roofline"] = {"bound": "hbm", "kernel": dom, "launches_per_step": n_dom, "achieved": ach, "peak": peak,
                           "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "traffic_source": traffic_src,
                           "peak_source": peak_src,
                           "algorithmic_bytes_per_launch": b_dom / n_dom, "avg_launch_ms": t_dom / n_dom,
                           "share_of_step": {g: v[0] / sum(x[0] for x in gsum.values()) for g, v in gsum.items()}}
        out["extra"].update({
            "fwd_applies_per_s_gather": 1000.0 / t_fwd_g, "adj_applies_per_s_gather": 1000.0 / t_adj_g,
            "fwd_hbm_frac": 2 * d * N * s / (t_fwd * 1e-3) / 1e9 / peak,
            "adj_hbm_frac": 2 * d * N * s / (t_adj * 1e-3) / 1e9 / peak,
            "kernel_ms": {**{k: v for k, v in passes.items() if v}, "CgUpdate(+3 dots)": t_upd, "Axpby": t_axp},
            "kernel_hbm_frac": {**{k: kbytes[k] / (v * 1e-3) / 1e9 / peak for k, v in passes.items() if v},
                                "CgUpdate(+3 dots)": 7 * N * s / (t_upd * 1e-3) / 1e9 / peak,
                                "Axpby": 3 * N * s / (t_axp * 1e-3) / 1e9 / peak}})

        # ---- CPU baseline on a bounded sample of the same workload
        if not args.no_cpu_baseline:
            csubs = cpu_sample_subs(subs, args.cpu_shape)
            cshape = tuple(n for sh, _ in csubs for n in sh)
            v, cores, t = cpu_cg_iters_per_s(csubs, offset, lik, args.dtype, iters=2)
            same = cshape == shape
            out["cpu_baseline"] = {"value": v * scale_to(cshape, shape), "unit": UNIT, "cores": cores, "kind": "port",
                                   "sample": f"oracle CG (the reference's scipy-pocketfft branch restated, workers={cores}) at "
                                             f"{'x'.join(map(str, cshape))}: 2 iterations, {v:.4f} it/s measured over {t:.1f}s"
                                             + ("" if same else ", scaled by N log N to the workload's grid"),
                                   "measured_value_at_sample": v}
        emit(out)
    del b
    if world > 1:
        cf.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def run_reference(args):
    """The reference's own CPU implementation of the path on the host cores.  /root/reference does not exist on
    the GPU box and its native arithmetic (ducc0) is not installable: this is the oracle port = the reference's
    `except ImportError` branch (scipy pocketfft) restated, with every host thread.  It is timed on the
    WORKLOAD'S OWN GRID when the host has the memory (1024^3 fp32 needs ~60 GB), one CG iteration per timed
    step, at most 2 timed steps so that the run ends within a few minutes; otherwise on the largest
    power-of-two sample that fits, scaled by N log N (stated in `sample`)."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    subs, offset, lik, desc, shape = resolve_config(args)
    N = float(np.prod(shape))
    s = 4 if args.dtype == "f32" else 8
    need_gb = 14.0 * N * s / 2 ** 30 + 4       # grid index (int64 x2), complex transform, temporaries
    mem = host_memory_gb()
    csubs = subs if (mem >= need_gb and not args.cpu_shape) else cpu_sample_subs(subs, args.cpu_shape)
    cshape = tuple(n for sh, _ in csubs for n in sh)
    iters = max(1, min(args.steps, 2))
    v, cores, t = cpu_cg_iters_per_s(csubs, offset, lik, args.dtype, iters=iters)
    same = cshape == shape
    val = v * scale_to(cshape, shape)
    out = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": int(os.environ.get("WORLD_SIZE", 1)),
           "steps": iters, "warmup": 0, "requested_steps": args.steps, "requested_warmup": args.warmup,
           "ms_per_step": 1000.0 / val, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
           "config": {"workload": desc + "; 1 step = 1 CG iteration = CF fwd + CF adj + vector updates",
                      "name": args.config, "likelihood": lik, "same_grid_as_gpu_arm": same,
                      "timed": f"{iters} CG iteration(s) at {'x'.join(map(str, cshape))} in {t:.1f} s on {cores} host threads"},
           "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                            "sample": f"oracle restatement of the reference CG + CF metric (ducc0 absent -> the reference's "
                                      f"own scipy pocketfft branch), {iters} iteration(s) at {'x'.join(map(str, cshape))} in {t:.1f}s"
                                      + ("" if same else f"; host has {mem:.0f} GB available, the full grid needs ~{need_gb:.0f} GB: "
                                                         "scaled by N log N")},
           "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0}
    emit(out)


_REAL_STDOUT = None


def guard_stdout():
    """Route fd 1 to stderr for the life of the process so that native libraries (NCCL prints its
    version banner on stdout) cannot add lines next to the ONE JSON line; emit() writes that line
    to the real stdout."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(out):
    line = (json.dumps(out) + "\n").encode()
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        os.write(1, line)
    else:
        os.write(_REAL_STDOUT, line)


if __name__ == "__main__":
    a = parse_args()
    guard_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
