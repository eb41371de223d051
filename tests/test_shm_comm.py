"""Shared-memory all-reduce of the host runtime (csrc/cpu/shm_comm.cpp) behind comm.inference_all_reduce.
Reference: csrc/cpu/comm/shm.cpp, comm/torch.py:186."""
import os

import pytest
import torch

from common import run_distributed


def _shm(world, dtype_name):
    import deepspeed_b200.comm as dist
    from deepspeed_b200.comm import shm
    os.environ["DSB200_SHM"] = "1"
    dtype = getattr(torch, dtype_name)
    rank = dist.get_rank()
    out = []
    for it, n in enumerate([1, 7, 1000, 4097, (9 << 20) + 3]):  # the last one needs several chunks of the 32 MiB buffer
        g = torch.Generator().manual_seed(100 * it + rank)
        x = torch.randn(n, generator=g).to(dtype)
        want = torch.zeros(n, dtype=torch.float32)
        for r in range(world):
            want += torch.randn(n, generator=torch.Generator().manual_seed(100 * it + r)).to(dtype).float()
        dist.inference_all_reduce(x)
        assert shm._STATE["handle"] is not None, "shared-memory path not taken"
        tol = 1e-5 if dtype == torch.float32 else 4e-2
        out.append(float((x.float() - want).abs().max() / want.abs().max().clamp(min=1.0)) < tol)
    # GPU-less fallbacks: non-contiguous tensors and other ops take the backend path and still reduce correctly
    y = torch.ones(6, 4)[:, ::2]
    dist.inference_all_reduce(y)
    out.append(bool((y == world).all()))
    shm.shutdown()
    return out


@pytest.mark.parametrize("dtype_name", ["float32", "bfloat16", "float16"])
def test_shm_all_reduce_matches_sum(dtype_name):
    world = 4
    res = run_distributed(_shm, world, args=(world, dtype_name))
    for r in range(world):
        assert all(res[r]), res
