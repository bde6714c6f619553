"""System diagnostics (reference: ``scaletorch/utils/env_utils.py:61-131``)."""
from __future__ import annotations

import platform
import socket

import torch

from .device import get_device_type


def get_system_info() -> dict:
    """Host / software / GPU diagnostics as a dict: hostname, platform, python, torch and CUDA versions, device count, name, memory and SM count."""
    info = {
        "hostname": socket.gethostname(),
        "platform": platform.platform(),
        "python": platform.python_version(),
        "torch": torch.__version__,
        "cuda": torch.version.cuda,
        "device_type": get_device_type(),
    }
    if get_device_type() == "cuda":
        info["device_count"] = torch.cuda.device_count()
        props = torch.cuda.get_device_properties(0)
        info["device_name"] = props.name
        info["device_memory_gb"] = round(props.total_memory / 1e9, 1)
        info["sm_count"] = props.multi_processor_count
        info["compute_capability"] = f"{props.major}.{props.minor}"
        try:
            info["nccl"] = ".".join(map(str, torch.cuda.nccl.version()))
        except Exception:
            pass
    try:
        from .. import _native

        info["native_library"] = str(_native.lib_path()) if _native.load() is not None else None
    except Exception:
        info["native_library"] = None
    return info


def init_dist_pytorch(backend: str | None = None, **kwargs) -> None:
    """Deprecated alias of :func:`scaletorch_b200.dist.init_dist` (kept for reference parity)."""
    from ..dist import init_dist

    init_dist("pytorch", backend=backend, **kwargs)


def cleanup_dist() -> None:
    """Deprecated alias of ``scaletorch_b200.dist.cleanup_dist`` kept for the reference's import path."""
    from ..dist import cleanup_dist as _c

    _c()
