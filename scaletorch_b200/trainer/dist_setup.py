"""Distributed start-up (reference: ``scaletorch/trainer/dist_setup.py:29-134``)."""
from __future__ import annotations

import datetime
import os

import torch
import torch.distributed as dist

from ..parallel.process_group import process_group_manager as pgm
from ..parallel.process_group import setup_process_group_manager
from ..utils.device import get_current_device, is_bf16_supported, set_device
from ..utils.logger_utils import get_logger

logger = get_logger(__name__)
DIST_TIMEOUT_S = 180


def get_dtype(config) -> torch.dtype:
    """bf16 when the device supports it and the config asks for it, else fp32.  (The reference ignores
    ``--dtype``; here ``--dtype float32`` is honoured.)"""
    want = getattr(config, "dtype", "bfloat16")
    if want in ("float32", "fp32"):
        return torch.float32
    if getattr(config, "use_cpu", False):
        return torch.float32
    return torch.bfloat16 if is_bf16_supported() else torch.float32


def initialize_distributed_training(config) -> tuple[int, int, int, torch.device]:
    """env ranks → ``init_process_group`` → ``set_device`` → ``ProcessGroupManager``.

    Returns ``(local_rank, global_rank, world_size, device)``.  With ``world_size == 1`` no process group
    is created and ``pgm`` stays falsy (reference ``dist_setup.py:120-122``).
    """
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    global_rank = int(os.environ.get("RANK", 0))
    world_size = int(os.environ.get("WORLD_SIZE", 1))
    use_cpu = bool(getattr(config, "use_cpu", False)) or not torch.cuda.is_available()
    device = get_current_device(use_cpu)
    if not use_cpu:
        set_device(device)
    # zig-zag context parallelism is a job-wide switch read by the loader, the RoPE positions and the ring attention
    from ..parallel.context_parallel import zigzag

    cp = getattr(config, "context_parallel_size", 1)
    seq = getattr(config, "sequence_length", None) or 0
    want_zz = bool(getattr(config, "cp_zigzag", False)) and cp > 1
    zigzag.set_enabled(want_zz and seq % (2 * cp) == 0)
    if want_zz and not zigzag.enabled() and int(os.environ.get("RANK", 0)) == 0:
        print(f"[scaletorch_b200] sequence_length {seq} is not divisible by 2*cp ({2 * cp}): contiguous context-parallel "
              f"split instead of zig-zag", flush=True)
    if world_size == 1:
        return local_rank, global_rank, world_size, device
    backend = "gloo" if use_cpu else getattr(config, "backend", "nccl")
    if not dist.is_initialized():
        kwargs = dict(backend=backend, init_method="env://", rank=global_rank, world_size=world_size,
                      timeout=datetime.timedelta(seconds=DIST_TIMEOUT_S))
        if backend == "nccl":
            kwargs["device_id"] = device
        dist.init_process_group(**kwargs)
    setup_process_group_manager(
        tp_size=config.tensor_parallel_size, cp_size=config.context_parallel_size,
        pp_size=config.pipeline_parallel_size, dp_size=config.data_parallel_size,
        ep_size=config.expert_parallel_size, ep_replicate_data=getattr(config, "ep_replicate_data", False))
    return local_rank, global_rank, world_size, device


def cleanup_distributed() -> None:
    """Tear down in the order CUDA graphs require: release captured steps, destroy the parallel groups, then the default process group."""
    from .graphed import release_all

    release_all()                                   # captured steps first: see graphed.release_all
    if pgm:
        try:
            pgm.cleanup()
        except Exception:
            pass
        pgm._instance = None
    if dist.is_initialized():
        dist.destroy_process_group()


# the reference's name (scaletorch/trainer/__init__.py:16, dist_setup.py)
cleanup_distributed_training = cleanup_distributed
