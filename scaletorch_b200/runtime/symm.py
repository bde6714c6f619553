"""Symmetric-memory runtime: per process-group peer-mapped buffers + signal pads.

One window per (group, purpose) is allocated through ``torch.distributed._symmetric_memory`` (CUDA VMM
allocation + handle exchange over the group's store); our kernels receive the raw peer pointers
(``buffer_ptrs``), the NVLS multicast pointer when the driver provides one, and the signal pads.
Nothing here launches kernels -- see ``runtime/fused_ops.py``.
"""
from __future__ import annotations

import os
from dataclasses import dataclass, field
from typing import Any

import torch
import torch.distributed as dist

from .. import _native as N

_windows: dict[tuple, "Window"] = {}
_ready_groups: dict[int, bool] = {}


@dataclass
class Window:
    tensor: torch.Tensor            # local view of the symmetric buffer (uint8)
    handle: Any                     # torch symmetric-memory handle
    rank: int
    world: int
    peer_ptrs: torch.Tensor         # int64 [world] device tensor of peer base pointers
    signal_ptrs: torch.Tensor       # int64 [world] device tensor of peer signal-pad pointers
    multicast_ptr: int = 0
    nbytes: int = 0
    epoch: int = 0                  # monotonically increasing flag value (avoids re-zeroing pads)
    extra: dict = field(default_factory=dict)


def has_kernel(name: str) -> bool:
    lib = N.load()
    return lib is not None and hasattr(lib, name)


def enabled() -> bool:
    return os.environ.get("SCALETORCH_B200_FUSED_COMM", "1") == "1" and torch.cuda.is_available()


def group_ready(group) -> bool:
    """True when ``group`` is an NCCL group whose ranks can all map each other's memory."""
    if group is None or not enabled() or not dist.is_initialized():
        return False
    key = id(group)
    if key not in _ready_groups:
        ok = False
        try:
            if dist.get_backend(group) == "nccl" and dist.get_world_size(group) > 1:
                import torch.distributed._symmetric_memory as sm  # noqa: F401

                ok = True
        except Exception:
            ok = False
        _ready_groups[key] = ok
    return _ready_groups[key]


def get_window(group, purpose: str, nbytes: int) -> Window:
    """Returns (allocating / growing collectively if needed) the symmetric window for ``purpose``.
    Must be called by every rank of ``group`` with the same arguments."""
    import torch.distributed._symmetric_memory as sm

    key = (id(group), purpose)
    w = _windows.get(key)
    if w is not None and w.nbytes >= nbytes:
        return w
    nbytes = max(int(nbytes), 1 << 20)
    nbytes = (nbytes + (1 << 21) - 1) // (1 << 21) * (1 << 21)
    dev = torch.device("cuda", torch.cuda.current_device())
    t = sm.empty(nbytes, dtype=torch.uint8, device=dev)
    h = sm.rendezvous(t, group=group)
    world, rank = h.world_size, h.rank
    peer = torch.tensor([int(p) for p in h.buffer_ptrs], dtype=torch.int64, device=dev)
    sig = torch.tensor([int(p) for p in h.signal_pad_ptrs], dtype=torch.int64, device=dev)
    mc = 0
    try:
        mc = int(h.multicast_ptr or 0)
    except Exception:
        mc = 0
    w = Window(tensor=t, handle=h, rank=rank, world=world, peer_ptrs=peer, signal_ptrs=sig, multicast_ptr=mc,
               nbytes=nbytes)
    _windows[key] = w
    return w


def reset() -> None:
    _windows.clear()
    _ready_groups.clear()
