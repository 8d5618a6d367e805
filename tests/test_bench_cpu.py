"""CPU-only checks of bench.py's host logic: config presets, the CPU arm on a tiny sample, the traffic
lookup, and the pre-timing multi-GPU parity check (verify_slab) under gloo with the kernel emulator."""

import json
import os
import subprocess
import sys

import numpy as np

import conftest

ROOT = conftest.ROOT


def test_configs_and_cpu_arm():
    import bench

    class A:
        config, shape, cpu_shape = "c5", "", ""
    subs, offset, lik, desc, shape = bench.resolve_config(A)
    assert shape == (2048, 2048, 128) and len(subs) == 2 and lik == "gaussian"
    A.config, A.shape = "c3", "32,32,32"
    subs, offset, lik, desc, shape = bench.resolve_config(A)
    assert shape == (32, 32, 32)
    v, cores, t = bench.cpu_cg_iters_per_s(subs, offset, lik, "f32", iters=2)
    assert v > 0 and cores >= 1
    small = bench.cpu_sample_subs(bench.CONFIGS["c4"][0])
    assert np.prod(small[0][0]) <= 2 ** 27 and small[0][0] == (512, 512, 512)
    small = bench.cpu_sample_subs(bench.CONFIGS["c5"][0])
    assert np.prod([n for sh, _ in small for n in sh]) <= 2 ** 27 and small[1][0] == (128,)
    # the product-space model runs through the CPU arm too
    v, _, _ = bench.cpu_cg_iters_per_s([((16, 16), bench.CF_KW_2D), ((8,), bench.CF_KW_FREQ)], bench.OFFSET_2D, "gaussian", "f64")
    assert v > 0


def test_reference_arm_prints_one_json_line():
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--config", "c3",
                          "--shape", "32,32,32", "--steps", "3", "--warmup", "1"], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [ln for ln in res.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    out = json.loads(lines[0])
    assert out["impl"] == "reference" and out["value"] > 0 and out["cpu_baseline"]["kind"] == "port"
    assert out["config"]["same_grid_as_gpu_arm"] and out["e2e"]["h2d_bytes_per_step"] == 0
    assert out["ms_per_step"] * out["steps"] < 600e3


def test_dram_traffic_lookup(tmp_path, monkeypatch):
    import bench
    monkeypatch.setattr(bench, "ROOT", str(tmp_path))
    assert bench.dram_traffic("c4", "f32", bench.G_ROWS)[0] is None
    os.makedirs(tmp_path / "profiles")
    db = {"c4/f32": {"source": "x.csv", "cases": {"rows+mul": {"dram_bytes": 10.0}, "rows+field+add+dot": {"dram_bytes": 14.0}}}}
    (tmp_path / "profiles" / "dram_bytes.json").write_text(json.dumps(db))
    assert bench.dram_traffic("c4", "f32", bench.G_ROWS) == (12.0, "x.csv")
    assert bench.dram_traffic("c4", "f32", bench.G_STRIDED3)[0] is None


def _verify_worker(rank, world, port, emu, result):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    from niftyb200 import _lib
    _lib.use_library_for_testing(emu, "cpu")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import bench
        import niftyb200 as nb
        err = bench.verify_slab(nb, torch, dist, rank, world, [((16, 16, 16), bench.CF_KW)])
        # product of two spaces (the c5 kind of grid) through the joint plan
        errp = bench.verify_slab(nb, torch, dist, rank, world, [((16, 8), bench.CF_KW_2D), ((16,), bench.CF_KW_FREQ)],
                                 bench.OFFSET_2D)
        if rank == 0:
            result["err"] = max(err, errp)
    finally:
        dist.destroy_process_group()


def test_verify_slab_gloo():
    import torch.multiprocessing as mp
    mgr = mp.Manager()
    result = mgr.dict()
    mp.spawn(_verify_worker, args=(2, 29581, conftest.build_emulator(), result), nprocs=2, join=True)
    assert result["err"] <= 1e-12
