"""world_size-2 test of the slab-decomposed path on CPU: gloo for the collectives, the kernel
emulator for the passes (tests/emu/README.md).  The N>1 GPU twin is test_gpu_distributed.py."""

import conftest
import dist_worker


def check(res, strict=True):
    tol = 1e-12 if res["dtype"] == "f64" else 1e-5
    for k in ("err_f", "err_a", "err_m", "err_c", "err_s"):
        assert res[k] <= tol, (k, res)
    if "nit" not in res:
        return
    assert res["err_static"] <= tol, res           # device-resident static_cg == host-driven static_cg
    if strict:
        assert res["nit"][0] == res["nit"][1] and res["nfev"][0] == res["nfev"][1], res
    else:   # a stopping iteration consistent with the reference's own sensitivity (parity_cases.acceptable_stops)
        assert res["nit"][0] in res["window"], \
            f"count {res['nit'][0]} (reference {res['nit'][1]}) not consistent with the reference: {res['window']}"
    assert res["info"][0] == res["info"][1], res
    if res["nit"][0] == res["nit"][1]:
        assert res["err_x"] <= (1e-8 if strict else 2e-2), res


def test_slab_two_ranks_gloo():
    res = dist_worker.launch(2, "gloo", conftest.build_emulator(), (16, 16, 16))
    check(res)


def test_slab_two_ranks_gloo_chunked_exchange(monkeypatch):
    # the exchange pipelined in 4 column chunks (what large grids use on the GPUs)
    monkeypatch.setenv("NIFTYB200_SLAB_CHUNKS", "4")
    res = dist_worker.launch(2, "gloo", conftest.build_emulator(), (16, 8, 32), port=29543)
    check(res)


def test_slab_two_ranks_gloo_long_rows():
    # rows of 512 scalars: the row pass with compile-time epilogues (EPI = 2 / 13) on the shifted-mirror
    # geometry of the slabs
    res = dist_worker.launch(2, "gloo", conftest.build_emulator(), (8, 8, 512), port=29545)
    check(res, strict=False)


def test_slab_four_ranks_gloo():
    res = dist_worker.launch(4, "gloo", conftest.build_emulator(), (16, 16, 8), port=29541)
    check(res)


def test_symmetric_split_is_a_partition():
    import numpy as np
    from niftyb200.distributed import SymmetricSplit
    for N, P in ((16, 1), (16, 2), (64, 4), (1024, 8)):
        s = SymmetricSplit(N, P)
        allk = np.concatenate([s.owned(r) for r in range(P)])
        assert sorted(allk.tolist()) == list(range(N))
        for r in range(P):
            own = s.owned(r)
            n = own.size
            for l in range(n):
                lm = (n - l - s.shift(r)) % n
                assert (N - own[l]) % N == own[lm]        # local mirror rule == global index negation


def test_slab_two_ranks_gloo_fp32():
    res = dist_worker.launch(2, "gloo", conftest.build_emulator(), (16, 16, 32), port=29547, mode="large", dtype="f32")
    check(res)


def test_sampled_metric_samples_sharded_over_ranks():
    res = dist_worker.launch_sampled(2, conftest.build_emulator(), (16, 16))
    assert res["err"] <= 1e-13 and res["errc"] <= 1e-13, res
    assert res["nit"][0] == res["nit"][1] and res["errx"] <= 1e-10, res


def test_slab_product_space_two_and_four_ranks_gloo():
    # (2-D space) x (1-D space) on the joint plan, slab-decomposed: only the row pass differs
    res = dist_worker.launch(2, "gloo", conftest.build_emulator(), (16, 8, 32), port=29549, mode="product")
    check(res)
    res = dist_worker.launch(4, "gloo", conftest.build_emulator(), (16, 16, 8), port=29551, mode="product", dtype="f32")
    check(res)
