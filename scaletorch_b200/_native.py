"""ctypes binding of ``libscaletorch_b200.so`` (see ``include/scaletorch_b200.h``).

The library is built in-tree by :mod:`scaletorch_b200._build`.  On a machine with
a CUDA device the ops in :mod:`scaletorch_b200.ops` *require* it (they raise if it
is missing instead of silently running a PyTorch fallback); on CPU-only machines
the pure-PyTorch oracle implementations are used.
"""
from __future__ import annotations

import ctypes
import os
import threading
from ctypes import c_char_p, c_float, c_int, c_long, c_void_p
from pathlib import Path

import torch

_LIB_PATH = Path(__file__).resolve().parent / "lib" / "libscaletorch_b200.so"
_lock = threading.Lock()
_lib: ctypes.CDLL | None = None
_load_error: str | None = None

# Number of kernels launched through this binding (bench.py reports it as gpu_launches), and the same per entry point.
launch_count = 0
launch_counts: dict[str, int] = {}

_P = c_void_p
_SIGS = {
    "stb_gemm_bf16": [_P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, _P],
    "stb_gemm_bf16_2cta": [_P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, _P],
    "stb_rmsnorm_fwd": [_P, _P, _P, _P, _P, _P, c_int, c_int, c_float, _P],
    "stb_rmsnorm_bwd": [_P, _P, _P, _P, _P, _P, _P, c_int, _P, c_int, c_int, _P],
    "stb_rmsnorm_bwd_noalloc": [_P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, _P],
    "stb_qknorm_rope_fwd": [_P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_float, _P],
    "stb_qknorm_rope_bwd": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, _P],
    "stb_flash_attn_fwd": [_P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_long, c_long, c_long, c_long,
                           c_long, c_long, c_long, c_long, c_long, c_float, c_int, _P],
    "stb_flash_attn_bwd": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int,
                           c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long,
                           c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_float, c_int, _P],
    "stb_flash_attn_fwd_block": [_P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int,
                                 c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_float, c_int, _P],
    "stb_flash_attn_bwd_block": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int,
                                 c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long,
                                 c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_float, c_int,
                                 c_int, c_int, c_int, c_int, _P],
    "stb_attn_bwd_delta": [_P, _P, _P, c_int, c_int, c_int, _P],
    "stb_flash_attn_bwd_fused": [_P, _P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int,
                                 c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long,
                                 c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_long, c_float, c_int, _P],
    "stb_qknorm_rope_bwd_f32q": [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_int,
                                 c_int, _P],
    "stb_swiglu_fwd": [_P, _P, c_long, c_int, _P],
    "stb_swiglu_bwd": [_P, _P, _P, c_long, c_int, _P],
    "stb_ce_stats": [_P, _P, _P, c_int, c_int, c_long, c_int, _P],
    "stb_ce_stats_chunk": [_P, _P, _P, c_int, c_int, c_long, c_int, c_int, _P],
    "stb_ce_finalize": [_P, _P, _P, _P, c_int, c_long, _P],
    "stb_ce_grad": [_P, _P, _P, c_int, c_int, c_long, c_int, c_float, _P, c_long, _P],
    "stb_adamw": [_P, c_int, _P, _P, _P, c_long, c_float, c_float, c_float, c_float, c_float, c_int, _P, c_float, _P],
    "stb_adamw_dyn": [_P, c_int, _P, _P, _P, c_long, c_float, c_float, c_float, c_float, _P, c_float, _P, _P],
    "stb_sumsq": [_P, c_long, _P, _P, _P],
    "stb_clip_coef": [_P, c_float, _P, _P, _P],
    "stb_cast_scale_f32_bf16": [_P, _P, c_long, c_float, _P],
    "stb_embedding_fwd": [_P, _P, _P, c_long, c_int, c_int, c_int, _P],
    "stb_embedding_bwd": [_P, _P, _P, c_long, c_int, c_int, c_int, _P],
}
_RESTYPES: dict = {}     # non-int return types of entry points registered by optional modules
_LONG_RET = {
    "stb_rmsnorm_bwd_scratch_elems": [c_int, c_int],
    "stb_sumsq_scratch_bytes": [],
}


def lib_path() -> Path:
    """Where the in-tree library is expected (``scaletorch_b200/lib/libscaletorch_b200.so``)."""
    return _LIB_PATH


def _bind(lib: ctypes.CDLL) -> None:
    lib.stb_last_error.restype = c_char_p
    lib.stb_last_error.argtypes = []
    lib.stb_abi_version.restype = c_int
    for name, sig in _SIGS.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.restype = _RESTYPES.get(name, c_int)
            fn.argtypes = sig
    for name, sig in _LONG_RET.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.restype = c_long
            fn.argtypes = sig


def register(name: str, argtypes: list, restype=c_int) -> None:
    """Registers the signature of an additional entry point (used by optional modules)."""
    _SIGS[name] = argtypes
    _RESTYPES[name] = restype
    if _lib is not None and hasattr(_lib, name):
        fn = getattr(_lib, name)
        fn.restype = restype
        fn.argtypes = argtypes


def load(required: bool = False) -> ctypes.CDLL | None:
    """Loads the native library (building it if the sources are newer is *not* done here)."""
    global _lib, _load_error
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not _LIB_PATH.exists():
            _load_error = f"{_LIB_PATH} does not exist (run `python -m scaletorch_b200._build`)"
        else:
            try:
                lib = ctypes.CDLL(str(_LIB_PATH), mode=ctypes.RTLD_GLOBAL)
                _bind(lib)
                _lib = lib
            except OSError as e:  # pragma: no cover - depends on the machine
                _load_error = f"cannot load {_LIB_PATH}: {e}"
    if _lib is None and required:
        raise RuntimeError(f"scaletorch_b200 native library unavailable: {_load_error}")
    return _lib


def available() -> bool:
    """True when CUDA kernels can actually run: library loaded and a CUDA device present."""
    return torch.cuda.is_available() and load() is not None


def use_native(t: torch.Tensor) -> bool:
    """Kernel dispatch rule.  CUDA tensors *must* take the native path."""
    if not t.is_cuda:
        return False
    if os.environ.get("SCALETORCH_B200_FORCE_TORCH", "0") == "1":
        return False
    load(required=True)
    return True


def stream_ptr() -> int:
    """The current CUDA stream as an integer handle for the C ABI."""
    return torch.cuda.current_stream().cuda_stream


def ptr(t: torch.Tensor | None) -> int | None:
    """``data_ptr()`` of a tensor for the C ABI (``None`` stays ``None`` = null pointer)."""
    return None if t is None else t.data_ptr()


def call(name: str, *args) -> None:
    """Calls ``name`` and raises ``RuntimeError`` with the library's message on failure."""
    global launch_count
    lib = load(required=True)
    rc = getattr(lib, name)(*args)
    if rc != 0:
        msg = lib.stb_last_error()
        raise RuntimeError(f"{name} failed ({rc}): {msg.decode() if msg else '?'}")
    launch_count += 1
    launch_counts[name] = launch_counts.get(name, 0) + 1


def call_long(name: str, *args) -> int:
    """Call an entry point and return its integer result instead of treating it as a status code."""
    lib = load(required=True)
    return int(getattr(lib, name)(*args))
