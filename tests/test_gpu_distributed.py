"""N>1 GPU twin of test_distributed_cpu.py on the GPUs of one box: the DEFAULT exchange of the slab path
(peer-memory stores through CUDA IPC, `nb2_strided_pass_scatter`) and NCCL for the scalars, against the
single-process oracle -- forward, adjoint, metric, curvature, five back-to-back transforms (double-buffer
rule) and a CG solve.  Skipped on boxes with fewer GPUs than ranks."""

import pytest
import torch

import dist_worker
from test_distributed_cpu import check

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world,shape", [(2, (64, 64, 64)), (4, (64, 128, 32)), (8, (128, 128, 64)), (8, (256, 256, 128))])
def test_slab_peer_exchange(world, shape):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    res = dist_worker.launch(world, "nccl", None, shape, port=29550 + world)
    assert res["p2p"] and res["world"] == world
    check(res, strict=False)


@pytest.mark.parametrize("world,shape,dtype", [(8, (512, 512, 512), "f32"), (4, (256, 256, 256), "f32"), (8, (256, 256, 256), "f64")])
def test_slab_peer_exchange_large(world, shape, dtype):
    """BASELINE-sized grid: forward / adjoint / metric / back-to-back transforms against the full oracle."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    res = dist_worker.launch(world, "nccl", None, shape, port=29560 + world, mode="large", dtype=dtype)
    assert res["p2p"] and res["world"] == world
    check(res)


def test_slab_nccl_all_to_all(monkeypatch):
    """The alternative data plane (NIFTYB200_SLAB_P2P=0: NCCL all_to_all_single per column chunk)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    monkeypatch.setenv("NIFTYB200_SLAB_P2P", "0")
    world = min(4, torch.cuda.device_count())
    res = dist_worker.launch(world, "nccl", None, (64, 128, 64), port=29571)
    assert not res["p2p"]
    check(res, strict=False)


@pytest.mark.parametrize("world,shape,dtype", [(2, (256, 128, 64), "f64"), (4, (512, 512, 32), "f32"), (8, (256, 256, 64), "f64")])
def test_slab_product_space(world, shape, dtype):
    """(2-D space) x (1-D frequency space) on the joint plan, slab-decomposed over the first space axis: the
    BASELINE configs[4] kind of grid on several GPUs (forward / adjoint / metric / back-to-back transforms)."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    res = dist_worker.launch(world, "nccl", None, shape, port=29580 + world, mode="product", dtype=dtype)
    assert res["p2p"] and res["world"] == world
    check(res)
