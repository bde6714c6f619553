"""Shared model utilities: attention-backend registry, RMSNorm module, RoPE helpers, weight init.

Reference: ``scaletorch/models/attention_utils.py:33-271``.  Back-ends operate on the token-major
``[B, S, H, D]`` layout and receive un-expanded GQA K/V (the reference uses ``[B, H, S, D]`` with K/V
repeated to the query head count).
"""
from __future__ import annotations

import math
from typing import Callable

import torch
import torch.nn as nn

from .. import ops
from ..ops.norm import apply_rotary_pos_emb, get_cos_sin, rotate_half  # noqa: F401  (re-exported)

AttentionFn = Callable[..., torch.Tensor]
_ATTENTION_BACKENDS: dict[str, AttentionFn] = {}


def register_attention_backend(name: str, fn: AttentionFn | None = None):
    """Registers ``fn(q, k, v, causal=True, scale=None) -> out`` under ``name`` (usable as a decorator)."""

    def _reg(f: AttentionFn) -> AttentionFn:
        _ATTENTION_BACKENDS[name] = f
        return f

    return _reg(fn) if fn is not None else _reg


def get_attention_backend(name: str) -> AttentionFn:
    """The attention function registered as ``name`` (``ring``, ``flash``, ``sdpa`` or a plug-in); ``KeyError`` lists the known names."""
    if name not in _ATTENTION_BACKENDS:
        raise KeyError(f"unknown attention backend '{name}'; registered: {sorted(_ATTENTION_BACKENDS)}")
    return _ATTENTION_BACKENDS[name]


def list_attention_backends() -> list[str]:
    """Names of the registered attention back-ends, sorted."""
    return sorted(_ATTENTION_BACKENDS)


def resolve_attention_backend_name(use_cp: bool, use_flash: bool) -> str:
    """ring > flash > sdpa (reference ``attention_utils.py:56-64``)."""
    if use_cp:
        return "ring"
    return "flash" if use_flash else "sdpa"


def flash_attention(q, k, v, causal: bool = True, scale: float | None = None) -> torch.Tensor:
    """The framework's flash kernel (tcgen05, native GQA) -- ``[B,S,H,D]`` in, ``[B,S,H,D]`` out."""
    return ops.flash_attention(q, k, v, causal, scale)[0]


def sdpa_attention(q, k, v, causal: bool = True, scale: float | None = None) -> torch.Tensor:
    """Eager fp32 oracle attention (the reference's "sdpa" name; used for debugging / tests)."""
    return ops.attention_ref(q, k, v, causal, scale)[0]


def _ring_attention(q, k, v, causal: bool = True, scale: float | None = None) -> torch.Tensor:
    from ..parallel.context_parallel.context_parallel import ring_attention

    return ring_attention(q, k, v, scale if scale is not None else 1.0 / math.sqrt(q.shape[-1]), causal)


register_attention_backend("flash", flash_attention)
register_attention_backend("sdpa", sdpa_attention)
register_attention_backend("ring", _ring_attention)


def _init_weights(w: torch.Tensor) -> None:
    """uniform(±1/sqrt(fan_in)) (reference ``attention_utils.py:160-167``)."""
    fan_in = w.shape[-1] if w.dim() > 1 else w.shape[0]
    bound = 1.0 / math.sqrt(fan_in)
    nn.init.uniform_(w, -bound, bound)


class RMSNorm(nn.Module):
    """``w * (x * rsqrt(mean(x²) + eps))`` with fp32 statistics (fused sm_100a kernel on CUDA)."""

    def __init__(self, hidden_size: int, eps: float = 1e-6):
        super().__init__()
        self.hidden_size = hidden_size
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(hidden_size))

    def reset_parameters(self) -> None:
        if self.weight.device.type != "meta":
            nn.init.ones_(self.weight)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.rmsnorm(x, self.weight, self.eps)

    def extra_repr(self) -> str:
        return f"{self.hidden_size}, eps={self.eps}"
