"""Availability checks for the symmetric-memory fused collectives (the windows themselves live in
``runtime/fused_ops.py``)."""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from .. import _native as N

_ready_groups: dict[int, bool] = {}
DEFAULT_WINDOW_BYTES = int(os.environ.get("SCALETORCH_B200_SYMM_WINDOW_MB", "1024")) << 20


def has_kernel(name: str) -> bool:
    """True when the in-tree library is loaded and exports ``name``."""
    lib = N.load()
    return lib is not None and hasattr(lib, name)


def enabled() -> bool:
    """Whether the peer-memory paths may be used at all: CUDA present and ``SCALETORCH_B200_FUSED_COMM`` not 0."""
    return os.environ.get("SCALETORCH_B200_FUSED_COMM", "1") == "1" and torch.cuda.is_available()


def group_ready(group) -> bool:
    """True when ``group`` is a multi-rank NCCL group and symmetric memory can be set up for it."""
    if group is None or not enabled() or not dist.is_initialized():
        return False
    key = id(group)
    if key not in _ready_groups:
        ok = False
        try:
            if dist.get_backend(group) == "nccl" and dist.get_world_size(group) > 1:
                import torch.distributed._symmetric_memory as sm  # noqa: F401

                from . import fused_ops

                fused_ops.get_group(group, DEFAULT_WINDOW_BYTES)   # collective: every rank of the group gets here
                ok = True
        except Exception as e:  # pragma: no cover - depends on the fabric
            import warnings

            warnings.warn(f"symmetric memory unavailable for this group, falling back to NCCL: {e}")
            ok = False
        _ready_groups[key] = ok
    return _ready_groups[key]


def region_bytes(group) -> int:
    """Size in bytes of one region (all-gather / reduce-scatter / all-reduce / spare) of the group's symmetric window."""
    from . import fused_ops

    return fused_ops.get_group(group, DEFAULT_WINDOW_BYTES).region_bytes


def reset() -> None:
    """Drop the readiness cache and the symmetric groups (before the process groups are destroyed)."""
    from . import fused_ops

    _ready_groups.clear()
    fused_ops.reset()
