"""Kernels written after the round's GPU budget was spent: compiled, never run.  Skipped unless
SCALETORCH_B200_EXPERIMENTAL=1 so an unvalidated kernel can never turn the suite red; first thing to run next round:

    SCALETORCH_B200_EXPERIMENTAL=1 timeout 120 python -m pytest tests/test_experimental_gpu.py -x -q -m gpu
"""
import os

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(os.environ.get("SCALETORCH_B200_EXPERIMENTAL", "0") != "1",
                                                  reason="experimental kernels: set SCALETORCH_B200_EXPERIMENTAL=1")]


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def _small_ar_worker(rank, world, port):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from scaletorch_b200.runtime.fused_ops import SmallAllReduce

        ar = SmallAllReduce.for_group(dist.group.WORLD)
        for n in (1, 3, 4, 1000, 4097):
            for _ in range(2):
                t = torch.randn(n, device="cuda")
                ref = t.clone()
                dist.all_reduce(ref)
                ar.all_reduce_(t)
                torch.cuda.synchronize()
                assert torch.allclose(t, ref, atol=1e-5), (n, (t - ref).abs().max())
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_small_all_reduce_two_gpus():
    import socket

    import torch.multiprocessing as mp

    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_small_ar_worker, args=(2, port), nprocs=2, join=True)
