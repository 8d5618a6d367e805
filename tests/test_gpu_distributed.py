"""N>1 GPU twin of test_distributed_cpu.py: NCCL all-to-all + CUDA passes on 2 (and 4) GPUs of
one box.  Skipped on boxes with a single GPU."""

import pytest
import torch

import dist_worker
from test_distributed_cpu import check

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world,shape", [(2, (64, 64, 64)), (4, (64, 128, 32)), (8, (128, 128, 64))])
def test_slab_nccl(world, shape):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    res = dist_worker.launch(world, "nccl", None, shape, port=29550 + world)
    check(res, strict=False)
