#!/usr/bin/env python3
"""bench.py -- metric-CG iterations/s (each = 1 CF forward + 1 CF adjoint apply + fused CG
vector updates) on a synthetic correlated field, BASELINE.json's metric and config.

  python bench.py --gpus N --steps K --warmup W [--dtype f32|f64] [--shape 1024,1024,1024]
  python bench.py --impl reference ...     # the reference's CPU algorithm on the host cores

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

import numpy as np  # noqa: E402

FUSED = "StridedPass(first axis)+amplitude field+direction update (fused axpby)"
METRIC = "metric_cg_iters_per_s"
UNIT = "CG iterations/s"
# demos/1_tomography.py:21-34 hyper-parameters
CF_KW = dict(fluctuations=(1.0, 0.5), loglogavgslope=(-5.0, 0.2), flexibility=(1.0, 0.2), asperity=(0.5, 5e-2))
OFFSET = (2.0, (1e-1, 3e-2))


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--shape", default="1024,1024,1024")
    ap.add_argument("--cpu-shape", default="512,512,512", help="bounded sample for the CPU baseline")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


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
def cpu_cg_iters_per_s(shape, dtype, iters=2):
    """Oracle (NumPy/SciPy restatement of the reference, oracle/) CG iterations on the host
    cores: reference fallback path scipy pocketfft with workers = all cores."""
    from oracle import correlated_field as ocf, metric as om, cg as ocg, transforms as otr
    cores = os.cpu_count() or 1
    otr.set_nthreads(cores)
    dt = np.float32 if dtype == "f32" else np.float64
    mk = ocf.CorrelatedFieldMaker("")
    mk.set_amplitude_total_offset(*OFFSET)
    mk.add_fluctuations(shape, 1.0 / shape[0], **CF_KW, non_parametric_kind="power")
    cf = mk.finalize()
    rng = np.random.default_rng(43)
    p = {k: rng.standard_normal(s) for k, s in cf.domain.items() if k != "xi"}
    amps, azm = cf.normalized_amplitudes(p), cf.azm(p)
    rng = np.random.default_rng(42)
    x0 = rng.standard_normal(shape, dtype=dt)
    w = np.exp(0.3 * rng.standard_normal(shape, dtype=dt))
    mat = om.cf_metric_xi(cf, [a.astype(dt) for a in amps], dt(azm), w)
    j = mat(x0) * dt(0.5) + x0
    t0 = time.perf_counter()
    res = ocg.cg_re(mat, j, x0, absdelta=0.0, miniter=iters, maxiter=iters, _raise_nonposdef=False)
    t = time.perf_counter() - t0
    # one iteration = one metric apply + vector updates; the solve did res.nfev applies
    return res.nfev / t, cores, t


def scale_to(shape_from, shape_to):
    """N log N work ratio used to express a bounded CPU sample in the headline workload's unit."""
    nf, nt = float(np.prod(shape_from)), float(np.prod(shape_to))
    return (nf * np.log2(nf)) / (nt * np.log2(nt))


# ---------------------------------------------------------------- GPU arm
def setup_problem(nb, torch, shape, tdt, seed=42):
    """Synthetic C3/C4 problem: xi ~ N(0,1), tomography hyper-parameters,
    Gaussian likelihood with a dense per-voxel inverse variance."""
    mk = nb.CorrelatedFieldMaker("", dtype=tdt)
    mk.set_amplitude_total_offset(*OFFSET)
    mk.add_fluctuations(shape, 1.0 / shape[0], **CF_KW, non_parametric_kind="power")
    cf = mk.finalize()
    rng = np.random.default_rng(43)
    p = {k: rng.standard_normal(s) for k, s in cf.domain.items() if not k.endswith("xi")}
    tab = cf.tables(p)
    gen = torch.Generator(device="cuda").manual_seed(seed)
    w = torch.exp(0.3 * torch.randn(shape, dtype=tdt, device="cuda", generator=gen))
    M = nb.CFMetric(cf, tab, w)
    x0 = torch.randn(shape, dtype=tdt, device="cuda", generator=gen)
    j = M(x0)
    j.mul_(0.5).add_(x0)
    return cf, tab, M, j, x0


def setup_slab_problem(nb, torch, shape, tdt, seed=42):
    """Same synthetic problem, slab-decomposed: local parts of w (layout Y) and x0 (layout X)."""
    mk = nb.CorrelatedFieldMaker("", dtype=tdt)
    mk.set_amplitude_total_offset(*OFFSET)
    mk.add_fluctuations(shape, 1.0 / shape[0], **CF_KW, non_parametric_kind="power")
    cf = nb.SlabCorrelatedField(mk.finalize())
    rng = np.random.default_rng(43)
    p = {k: rng.standard_normal(s) for k, s in cf.cf.domain.items() if not k.endswith("xi")}
    tab = cf.tables(p)
    gen = torch.Generator(device="cuda").manual_seed(seed)
    w = torch.exp(0.3 * torch.randn(cf.position_shape, dtype=tdt, device="cuda", generator=gen))
    M = nb.CFMetric(cf, tab, w)
    x0 = torch.randn(cf.latent_shape, dtype=tdt, device="cuda", generator=gen)
    j = M(x0)
    j.mul_(0.5).add_(x0)
    return cf, tab, M, j, x0


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
    shape = tuple(int(s) for s in args.shape.split(","))
    tdt = torch.float32 if args.dtype == "f32" else torch.float64
    s = 4 if args.dtype == "f32" else 8
    N, d = int(np.prod(shape)), len(shape)
    peak, peak_src = peaks()

    if world > 1:
        cf, tab, M, j, x0 = setup_slab_problem(nb, torch, shape, tdt, seed=42 + rank)
    else:
        cf, tab, M, j, x0 = setup_problem(nb, torch, shape, tdt, seed=42)
    loop = cgm.CGLoop(M, j, x0)
    loop.run(args.warmup)
    n0 = lib.nb2_launch_count()
    with ClockSampler(local) as clk:
        ms = time_region(torch, dist, lambda: loop.run(args.steps))
    launches = int(lib.nb2_launch_count() - n0)
    ms_step = ms / args.steps
    value = 1000.0 / ms_step                  # ONE global grid, slab-decomposed over the ranks when world > 1

    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if world > 1 else "weak",
           "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
           "config": {"workload": f"3D correlated field {'x'.join(map(str, shape))} metric-CG "
                                  "(configs[3] grid; 1 step = 1 CG iteration = CF fwd + CF adj + fused updates)",
                      "l2": "inputs larger than L2 (each vector %.2f GB per GPU)" % (N * s / 1e9 / world),
                      "multi_gpu": (f"slab-decomposed over {world} GPUs: one NCCL all-to-all per transform, "
                                    "all-reduce of <=3 fp64 scalars per dot") if world > 1 else "single GPU"},
           "gpu_launches": launches, "clocks": clk.summary()}
    B_cg = (4 * d + 13) * N * s                # SURVEY.md 8(d) nominal bytes per CG iteration
    out["extra"] = {"cg_nominal_bytes": B_cg, "cg_hbm_frac_of_measured": B_cg / world / (ms_step * 1e-3) / 1e9 / peak}
    if world > 1:
        # combined roofline: max(HBM time of the local share, NVLink time of two all-to-alls)
        nvl = 770.0   # GB/s per direction per GPU, measured peer copy (B200_PROFILING.md)
        t_hbm = B_cg / world / (peak * 1e9)
        t_nvl = 2 * (world - 1) / world * N * s / world / (nvl * 1e9)
        out["extra"].update({"combined_roofline_ms": max(t_hbm, t_nvl) * 1e3, "t_hbm_ms": t_hbm * 1e3, "t_nvlink_ms": t_nvl * 1e3,
                             "frac_of_combined_roofline": max(t_hbm, t_nvl) * 1e3 / ms_step, "nvlink_gbs_assumed": nvl})

    if rank == 0 and world > 1:
        ach = B_cg / world / (ms_step * 1e-3) / 1e9
        out["roofline"] = {"bound": "hbm", "kernel": "whole CG iteration, local share (9 kernels + 2 all-to-alls)",
                           "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                           "peak_source": peak_src, "algorithmic_bytes_per_launch": B_cg / world}
        emit(out)
    if rank == 0 and world == 1:
        # ---- per-kernel timing (CUDA events, same stream), applies/s, roofline of the dominant kernel
        plan = cf.plans[0]
        a, b = x0, torch.empty_like(x0)
        reps = max(3, min(10, args.steps))

        def tk(fn):
            fn()
            return time_region(torch, None, lambda: [fn() for _ in range(reps)]) / reps

        tabs = M.tab          # amplitude gather resolved once per expansion point (streamed field)
        t_fwd = tk(lambda: cf.apply_xi(a, tabs, out=b))
        t_adj = tk(lambda: cf.adjoint_xi(a, tabs, out=b))
        t_fwd_g = tk(lambda: cf.apply_xi(a, tab, out=b))
        t_adj_g = tk(lambda: cf.adjoint_xi(a, tab, out=b))
        names = {0: "StridedPass(axis0)", 1: "StridedPass(axis1)", 2: "RowPass(contiguous+Hartley combine)"}
        passes = {}
        if plan.is_fast:
            pro = PointOp(tab.amp_mode, tab.amp, tab.azm)
            for which in range(3):
                if len(shape) - 3 + which < 0:
                    continue
                passes[names[which]] = tk(lambda w=which: plan.run_pass(w, a, b))
            if len(shape) > 1:
                pros = PointOp(tabs.amp_mode, tabs.amp, tabs.azm, amp_sym=tabs.sym)
                passes["StridedPass(first axis)+amplitude field"] = tk(lambda: plan.run_pass(3 - len(shape), a, b, pro=pros))
                passes["StridedPass(first axis)+amplitude gather"] = tk(lambda: plan.run_pass(3 - len(shape), a, b, pro=pro))
                if getattr(M, "fused_direction_update", False):
                    # the first pass of the metric apply as the CG loop runs it: d <- r + beta d in place, x field
                    prod = PointOp(tabs.amp_mode, tabs.amp, tabs.azm, amp_sym=tabs.sym, pre_add=j, pre_scale=0.5, pre_out=a)
                    passes[FUSED] = tk(lambda: plan.run_pass(3 - len(shape), a, b, pro=prod))
            passes["RowPass+weight (metric fwd epilogue)"] = tk(lambda: plan.run_pass(
                2, a, b, epi=PointOp(scale=cf.dvol, mul=M.w)))
            passes["RowPass+amplitude field+add+dot (metric adj epilogue)"] = tk(lambda: plan.run_pass(
                2, a, b, epi=PointOp(tabs.amp_mode, tabs.amp, tabs.azm * cf.dvol, add=x0, dot=x0, dot_out=M._curv,
                                     amp_sym=tabs.sym)))
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
            groups["strided pass (StridedTma / StridedPass): first axis + amplitude field, axis 1, axis 0, axis 1"] = \
                [f0, names[1], names[0], names[1]]
        elif len(shape) == 2 and passes.get(f0) and passes.get(names[1]):
            groups["strided pass (StridedTma / StridedPass): axis 1 + amplitude field, axis 1"] = [f0, names[1]]
        rk = ["RowPass+weight (metric fwd epilogue)", "RowPass+amplitude field+add+dot (metric adj epilogue)"]
        if all(passes.get(k) for k in rk):
            groups["row pass (RowPass): + weight, + amplitude field + add + curvature dot"] = rk
        gsum = {g: (sum(passes[k] for k in ks), sum(kbytes[k] for k in ks), len(ks)) for g, ks in groups.items()}
        gsum["CgUpdate (+3 dots)"] = (t_upd, 7 * N * s, 1)
        if not fused:       # otherwise the direction update rides on the first strided pass
            gsum["Axpby"] = (t_axp, 3 * N * s, 1)
        dom = max(gsum, key=lambda g: gsum[g][0])
        t_dom, b_dom, n_dom = gsum[dom]
        ach = b_dom / (t_dom * 1e-3) / 1e9
        # DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum, one `ncu --set full` capture,
        # profiles/r1_ncu_full_v3_passes_f32_1024.csv), known for the headline grid only
        traffic = None
        if shape == (1024, 1024, 1024) and args.dtype == "f32" and dom.startswith("strided") and not fused:
            traffic = (10.701343e9 + 8.532932e9 + 8.541014e9 + 8.532932e9) / 4
        out["roofline"] = {"bound": "hbm", "kernel": dom, "launches_per_step": n_dom, "achieved": ach, "peak": peak,
                           "unit": "GB/s", "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                           "algorithmic_bytes_per_launch": b_dom / n_dom, "avg_launch_ms": t_dom / n_dom,
                           "share_of_step": {g: v[0] / sum(x[0] for x in gsum.values()) for g, v in gsum.items()}}
        out["extra"].update({
            "fwd_applies_per_s": 1000.0 / t_fwd, "adj_applies_per_s": 1000.0 / t_adj,
            "fwd_adj_pairs_per_s": 1000.0 / (t_fwd + t_adj),
            "fwd_applies_per_s_gather": 1000.0 / t_fwd_g, "adj_applies_per_s_gather": 1000.0 / t_adj_g,
            "fwd_hbm_frac": 2 * d * N * s / (t_fwd * 1e-3) / 1e9 / peak,
            "adj_hbm_frac": 2 * d * N * s / (t_adj * 1e-3) / 1e9 / peak,
            "kernel_ms": {**{k: v for k, v in passes.items() if v}, "CgUpdate(+3 dots)": t_upd, "Axpby": t_axp},
            "kernel_hbm_frac": {**{k: kbytes[k] / (v * 1e-3) / 1e9 / peak for k, v in passes.items() if v},
                                "CgUpdate(+3 dots)": 7 * N * s / (t_upd * 1e-3) / 1e9 / peak,
                                "Axpby": 3 * N * s / (t_axp * 1e-3) / 1e9 / peak}})
        del b

        # ---- end to end through the public API with HOST buffers
        if not args.no_e2e:
            jh = torch.empty(shape, dtype=tdt, pin_memory=True).copy_(j)
            xh = torch.empty(shape, dtype=tdt, pin_memory=True).copy_(x0)
            oh = torch.empty(shape, dtype=tdt, pin_memory=True)
            K = args.steps

            def solve():
                jd, xd = jh.to("cuda", non_blocking=True), xh.to("cuda", non_blocking=True)
                xs, info = nb.cg(M, jd, xd, absdelta=0.0, miniter=K, maxiter=K, _raise_nonposdef=False)
                oh.copy_(xs, non_blocking=True)
                torch.cuda.synchronize()
                return info

            solve()
            t0 = time.perf_counter()
            solve()
            te = time.perf_counter() - t0
            out["e2e"] = {"value": K / te, "unit": UNIT, "h2d_bytes_per_step": 2 * N * s / K,
                          "d2h_bytes_per_step": N * s / K + 24,
                          "what": f"nb.cg(...) solve of {K} iterations, j and x0 from pinned host memory, "
                                  "solution back to host, per-iteration scalars read by the host"}
            del jh, xh, oh

        # ---- CPU baseline on a bounded sample
        if not args.no_cpu_baseline and world == 1:
            cshape = tuple(int(v) for v in args.cpu_shape.split(","))
            cshape = cshape if np.prod(cshape) <= N else shape
            v, cores, t = cpu_cg_iters_per_s(cshape, args.dtype)
            out["cpu_baseline"] = {"value": v * scale_to(cshape, shape), "unit": UNIT, "cores": cores, "kind": "port",
                                   "sample": f"oracle CG (scipy pocketfft, workers={cores}) at {'x'.join(map(str, cshape))}: "
                                             f"{v:.4f} it/s measured over {t:.1f}s, scaled by N log N to the headline grid",
                                   "measured_value_at_sample": v}
        emit(out)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def run_reference(args):
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    shape = tuple(int(s) for s in args.shape.split(","))
    cshape = tuple(int(v) for v in args.cpu_shape.split(","))
    cshape = cshape if np.prod(cshape) <= np.prod(shape) else shape
    iters = max(1, min(args.steps, 3))
    v, cores, t = cpu_cg_iters_per_s(cshape, args.dtype, iters=iters)
    val = v * scale_to(cshape, shape)
    out = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": int(os.environ.get("WORLD_SIZE", 1)),
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 / val, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
           "config": {"workload": f"3D correlated field {'x'.join(map(str, shape))} metric-CG, CPU arm on a bounded "
                                  f"sample {'x'.join(map(str, cshape))} scaled by N log N"},
           "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                            "sample": f"oracle restatement of the reference CG + CF metric (ducc0 absent -> the reference's "
                                      f"own scipy pocketfft branch), {iters} iterations at {'x'.join(map(str, cshape))} in {t:.1f}s"},
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
