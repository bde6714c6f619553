"""Process-group bootstrap and rank queries (reference: ``scaletorch/dist/utils.py:26-620``).

B200-only: back-ends are ``nccl`` (CUDA) and ``gloo`` (CPU tests); the NPU/MLU/MUSA branches and the
deepspeed/colossalai init back-ends of the reference are dropped.  ``cast_data_device`` /
``get_comm_device`` are kept because gloo groups need CPU tensors.
"""
from __future__ import annotations

import datetime
import functools
import os
import subprocess
from collections.abc import Iterable, Mapping
from typing import Callable

import torch
import torch.distributed as torch_dist
from torch import Tensor
from torch.distributed import ProcessGroup

_LOCAL_PROCESS_GROUP: ProcessGroup | None = None


def is_distributed() -> bool:
    """True once ``torch.distributed`` is available and the default process group has been initialised."""
    return torch_dist.is_available() and torch_dist.is_initialized()


def get_local_group() -> ProcessGroup | None:
    """The process group of the ranks on this node (``init_local_group``); ``None`` outside a distributed run."""
    if not is_distributed():
        return None
    if _LOCAL_PROCESS_GROUP is None:
        raise RuntimeError("local process group is not created, call init_local_group first")
    return _LOCAL_PROCESS_GROUP


def get_default_group() -> ProcessGroup | None:
    """The world process group, or ``None`` outside a distributed run."""
    return torch_dist.distributed_c10d._get_default_group() if is_distributed() else None


def cleanup_dist() -> None:
    """Destroy the default process group if there is one (safe to call twice)."""
    if is_distributed():
        torch_dist.destroy_process_group()


def new_group(ranks, **kwargs):
    """``torch.distributed.new_group`` (collective over the world); ``None`` outside a distributed run."""
    return torch_dist.new_group(ranks=ranks, **kwargs) if is_distributed() else None


def destroy_group(group) -> None:
    """Destroy one sub-group; a no-op for ``None`` or outside a distributed run."""
    if is_distributed() and group is not None:
        torch_dist.destroy_process_group(group)


def infer_launcher() -> str:
    """Guess the launcher from the environment: ``pytorch`` (torchrun), ``slurm``, ``mpi`` or ``none``."""
    if "WORLD_SIZE" in os.environ:
        return "pytorch"
    if "SLURM_NTASKS" in os.environ:
        return "slurm"
    if "OMPI_COMM_WORLD_LOCAL_RANK" in os.environ:
        return "mpi"
    return "none"


def init_dist(launcher: str = "pytorch", backend: str | None = None, timeout_s: int = 180, **kwargs) -> None:
    """Initialises ``torch.distributed`` for the given launcher ('pytorch' | 'slurm' | 'mpi')."""
    if is_distributed():
        return
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    kwargs.setdefault("timeout", datetime.timedelta(seconds=timeout_s))
    if launcher == "pytorch":
        _init_dist_pytorch(backend, **kwargs)
    elif launcher == "mpi":
        _init_dist_mpi(backend, **kwargs)
    elif launcher == "slurm":
        _init_dist_slurm(backend, **kwargs)
    else:
        raise ValueError(f"Invalid launcher type: {launcher}")


def _set_cuda_device(local_rank: int) -> None:
    if torch.cuda.is_available():
        torch.cuda.set_device(local_rank % torch.cuda.device_count())


def _init_dist_pytorch(backend: str, **kwargs) -> None:
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if backend == "nccl":
        _set_cuda_device(local_rank)
        kwargs.setdefault("device_id", torch.device("cuda", torch.cuda.current_device()))
    torch_dist.init_process_group(backend=backend, **kwargs)


def _init_dist_mpi(backend: str, **kwargs) -> None:
    local_rank = int(os.environ["OMPI_COMM_WORLD_LOCAL_RANK"])
    os.environ.setdefault("MASTER_PORT", "29500")
    if "MASTER_ADDR" not in os.environ:
        raise KeyError("The environment variable MASTER_ADDR is not set")
    os.environ["WORLD_SIZE"] = os.environ["OMPI_COMM_WORLD_SIZE"]
    os.environ["RANK"] = os.environ["OMPI_COMM_WORLD_RANK"]
    os.environ["LOCAL_RANK"] = str(local_rank)
    if backend == "nccl":
        _set_cuda_device(local_rank)
    torch_dist.init_process_group(backend=backend, **kwargs)


def _init_dist_slurm(backend: str, port: int | None = None, **kwargs) -> None:
    proc_id = int(os.environ["SLURM_PROCID"])
    ntasks = int(os.environ["SLURM_NTASKS"])
    node_list = os.environ["SLURM_NODELIST"]
    addr = subprocess.getoutput(f"scontrol show hostname {node_list} | head -n1")
    if port is not None:
        os.environ["MASTER_PORT"] = str(port)
    os.environ.setdefault("MASTER_PORT", "29500")
    os.environ.setdefault("MASTER_ADDR", addr)
    os.environ["WORLD_SIZE"] = str(ntasks)
    local_rank = proc_id % max(torch.cuda.device_count(), 1)
    os.environ["LOCAL_RANK"] = str(local_rank)
    os.environ["RANK"] = str(proc_id)
    if backend == "nccl":
        _set_cuda_device(local_rank)
    torch_dist.init_process_group(backend=backend, **kwargs)


def init_local_group(node_rank: int, num_gpus_per_node: int) -> None:
    """Creates the intra-node group (ranks ``[node_rank*n, (node_rank+1)*n)``)."""
    global _LOCAL_PROCESS_GROUP
    assert _LOCAL_PROCESS_GROUP is None
    world = get_world_size()
    for n in range(world // num_gpus_per_node):
        ranks = list(range(n * num_gpus_per_node, (n + 1) * num_gpus_per_node))
        g = torch_dist.new_group(ranks)
        if n == node_rank:
            _LOCAL_PROCESS_GROUP = g


def get_backend(group: ProcessGroup | None = None) -> str | None:
    """Backend name of ``group`` (default: world), ``None`` outside a distributed run."""
    if not is_distributed():
        return None
    return torch_dist.get_backend(group or get_default_group())


def get_world_size(group: ProcessGroup | None = None) -> int:
    """Number of ranks in ``group`` (default: world); 1 outside a distributed run."""
    if not is_distributed():
        return 1
    return torch_dist.get_world_size(group or get_default_group())


def get_rank(group: ProcessGroup | None = None) -> int:
    """Rank of this process inside ``group`` (default: world); 0 outside a distributed run."""
    if not is_distributed():
        return 0
    return torch_dist.get_rank(group or get_default_group())


def get_local_size() -> int:
    """Number of ranks on this node; 1 outside a distributed run."""
    if not is_distributed():
        return 1
    return torch_dist.get_world_size(get_local_group())


def get_local_rank() -> int:
    """Rank of this process on its node (falls back to ``LOCAL_RANK`` when no local group was created)."""
    if not is_distributed():
        return 0
    if _LOCAL_PROCESS_GROUP is None:
        return int(os.environ.get("LOCAL_RANK", 0))
    return torch_dist.get_rank(_LOCAL_PROCESS_GROUP)


def get_dist_info(group: ProcessGroup | None = None) -> tuple[int, int]:
    """``(rank, world_size)`` of ``group``."""
    return get_rank(group), get_world_size(group)


def is_main_process(group: ProcessGroup | None = None) -> bool:
    """True on rank 0 of ``group``."""
    return get_rank(group) == 0


def master_only(func: Callable) -> Callable:
    """Decorator: run the function on rank 0 only, return ``None`` elsewhere."""
    @functools.wraps(func)
    def wrapper(*args, **kwargs):
        if is_main_process():
            return func(*args, **kwargs)
        return None

    return wrapper


def barrier(group: ProcessGroup | None = None) -> None:
    """Block until every rank of ``group`` arrives; a no-op outside a distributed run."""
    if is_distributed():
        torch_dist.barrier(group or get_default_group())


def _iter_data(data):
    if isinstance(data, Tensor):
        yield data
    elif isinstance(data, Mapping):
        for v in data.values():
            yield from _iter_data(v)
    elif isinstance(data, Iterable) and not isinstance(data, (str, bytes)):
        for v in data:
            yield from _iter_data(v)


def get_data_device(data) -> torch.device:
    """Device shared by every tensor in ``data`` (tensor / mapping / iterable)."""
    dev = None
    for t in _iter_data(data):
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise ValueError(f"device type in data should be consistent, but got {t.device} and {dev}")
    if dev is None:
        raise ValueError("data should be a Tensor, sequence of tensors or dict with at least one tensor")
    return dev


def get_comm_device(group: ProcessGroup | None = None) -> torch.device:
    """Device tensors must live on for ``group``'s backend (cuda for nccl, cpu for gloo)."""
    if get_backend(group) == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def cast_data_device(data, device: torch.device, out=None):
    """Moves (nested) tensors to ``device``; with ``out`` copies the moved values into it in place."""
    if isinstance(data, Tensor):
        moved = data if data.device == device else data.to(device)
        if out is not None:
            if not isinstance(out, Tensor):
                raise TypeError("out should be the same type as data")
            out.copy_(moved)
        return moved
    if isinstance(data, Mapping):
        if out is not None and (not isinstance(out, Mapping) or len(out) != len(data)):
            raise ValueError("out should be a mapping of the same length as data")
        res = {k: cast_data_device(v, device, None if out is None else out[k]) for k, v in data.items()}
        return type(data)(res) if not isinstance(data, dict) else res
    if isinstance(data, Iterable) and not isinstance(data, (str, bytes)):
        data = list(data)
        outs = list(out) if out is not None else [None] * len(data)
        if len(outs) != len(data):
            raise ValueError("out should be an iterable of the same length as data")
        return [cast_data_device(v, device, o) for v, o in zip(data, outs)]
    raise TypeError(f"data should be a Tensor, list of tensors or dict, but got {type(data)}")
