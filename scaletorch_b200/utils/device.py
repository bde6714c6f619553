"""Device helpers for the single supported accelerator family: NVIDIA B200 (CUDA), with CPU as the
test / gloo path.

Reference: ``scaletorch/utils/device.py:24-298`` -- the multi-vendor probing (npu/xpu/mlu/musa/mps)
is deliberately dropped (BASELINE north star); the public function names are kept so reference
users find what they expect.  The peak-FLOPS registry gains B200 entries (the reference has none
and falls back to 312e12 on CUDA, ``utils/device.py:288-296``).
"""
from __future__ import annotations

import functools
import json
import os

import torch

from ..env import ENV_DEVICE_FLOPS, ENV_LOCAL_RANK, ENV_RANK, ENV_WORLD_SIZE

# dense bf16 peak FLOP/s by device-name substring (first match wins)
_DEVICE_FLOPS: list[tuple[str, float]] = [
    ("B200", 2.25e15),
    ("B300", 2.25e15),
    ("GB200", 2.5e15),
    ("H200", 989e12),
    ("H100", 989e12),
    ("A100", 312e12),
]
_CUDA_FALLBACK_FLOPS = 312e12


@functools.lru_cache(maxsize=None)
def get_device_type() -> str:
    """``cuda`` when a GPU is visible, else ``cpu`` (cached; the reference's seven-way probe is reduced to the two targets built here)."""
    return "cuda" if torch.cuda.is_available() else "cpu"


def is_accelerator_available() -> bool:
    """True when a CUDA device is visible."""
    return get_device_type() == "cuda"


def get_dist_backend() -> str:
    """``nccl`` on CUDA, ``gloo`` on CPU."""
    return "nccl" if get_device_type() == "cuda" else "gloo"


def get_visible_devices_keyword() -> str:
    """Name of the environment variable that restricts visible devices (``CUDA_VISIBLE_DEVICES``)."""
    return "CUDA_VISIBLE_DEVICES"


def get_dist_info() -> tuple[int, int, int]:
    """(rank, world_size, local_rank) from the launcher environment."""
    return (int(os.environ.get(ENV_RANK, 0)), int(os.environ.get(ENV_WORLD_SIZE, 1)), int(os.environ.get(ENV_LOCAL_RANK, 0)))


def get_current_device(use_cpu: bool = False) -> torch.device:
    """This rank's device: ``cuda:<LOCAL_RANK mod device_count>``, or ``cpu`` when ``use_cpu`` or no GPU is visible."""
    if use_cpu or get_device_type() == "cpu":
        return torch.device("cpu")
    return torch.device("cuda", int(os.environ.get(ENV_LOCAL_RANK, 0)) % max(torch.cuda.device_count(), 1))


def set_device(device: torch.device | int) -> None:
    """Make ``device`` (a ``torch.device`` or a CUDA index) current; a no-op for CPU."""
    if isinstance(device, int):
        device = torch.device("cuda", device) if get_device_type() == "cuda" else torch.device("cpu")
    if device.type == "cuda":
        torch.cuda.set_device(device)


def get_device_count() -> int:
    """Number of visible CUDA devices (0 on a CPU-only machine)."""
    return torch.cuda.device_count() if get_device_type() == "cuda" else 0


def synchronize() -> None:
    """Wait for all queued work on the current CUDA device; a no-op on CPU."""
    if get_device_type() == "cuda":
        torch.cuda.synchronize()


def empty_cache() -> None:
    """Release the caching allocator's unused blocks; a no-op on CPU."""
    if get_device_type() == "cuda":
        torch.cuda.empty_cache()


def memory_reserved() -> int:
    """Bytes reserved by the caching allocator on the current device (0 on CPU)."""
    return torch.cuda.memory_reserved() if get_device_type() == "cuda" else 0


def memory_allocated() -> int:
    """Bytes currently allocated to tensors on the current device (0 on CPU)."""
    return torch.cuda.memory_allocated() if get_device_type() == "cuda" else 0


def max_memory_allocated() -> int:
    """Peak of ``memory_allocated`` since the last reset (0 on CPU)."""
    return torch.cuda.max_memory_allocated() if get_device_type() == "cuda" else 0


def reset_peak_memory_stats() -> None:
    """Restart the peak-memory statistics of the current device."""
    if get_device_type() == "cuda":
        torch.cuda.reset_peak_memory_stats()


def is_bf16_supported() -> bool:
    """Whether the compute dtype can be bf16 (always true on B200; true on CPU, where it is merely slow)."""
    if get_device_type() == "cuda":
        return torch.cuda.is_bf16_supported()
    return True  # CPU bf16 works (slowly); tests use fp32 explicitly


def manual_seed_all(seed: int) -> None:
    """Seed the CPU generator and every CUDA generator."""
    torch.manual_seed(seed)
    if get_device_type() == "cuda":
        torch.cuda.manual_seed_all(seed)


def register_device_flops(name_substring: str, flops: float) -> None:
    """Registers / overrides a peak-FLOPS entry (checked before the built-in table)."""
    if flops <= 0:
        raise ValueError("flops must be positive")
    _DEVICE_FLOPS.insert(0, (name_substring, float(flops)))


def get_measured_peaks() -> dict:
    """Driver-measured box peaks (``MEASURED_PEAKS.json`` at the repo root), or ``{}``."""
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return json.load(f)
    except Exception:
        return {}


def get_theoretical_flops() -> float:
    """Peak dense bf16 FLOP/s used as the MFU denominator.

    Order: ``SCALETORCH_DEVICE_FLOPS`` env override → registry by device name → CUDA fallback.
    """
    env = os.environ.get(ENV_DEVICE_FLOPS)
    if env:
        try:
            v = float(env)
            if v > 0:
                return v
        except ValueError:
            pass
    if get_device_type() == "cuda":
        name = torch.cuda.get_device_name(0)
        for sub, flops in _DEVICE_FLOPS:
            if sub.lower() in name.lower():
                return flops
        return _CUDA_FALLBACK_FLOPS
    return 1e12  # CPU placeholder so MFU stays finite in tests
