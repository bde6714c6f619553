"""SwiGLU on a fused ``gate|up`` projection.

The framework stores ``gate_proj`` and ``up_proj`` as one ``[2I, H]`` weight so a single
GEMM produces ``[T, 2I]`` (gate in columns ``[0, I)``, up in ``[I, 2I)``); this op computes
``silu(gate) * up``.  Reference: ``scaletorch/models/model_qwen3.py:85-87``,
``scaletorch/models/llama.py:247-249``.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .. import _native as N


def swiglu_ref(gu: torch.Tensor) -> torch.Tensor:
    """PyTorch oracle: ``silu(gate) * up`` for ``gu = [gate | up]`` along the last dimension."""
    g, u = gu.chunk(2, dim=-1)
    return F.silu(g) * u


class _SwiGLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gu):
        I = gu.shape[-1] // 2
        g2 = gu.reshape(-1, 2 * I)
        act = torch.empty((g2.shape[0], I), dtype=gu.dtype, device=gu.device)
        N.call("stb_swiglu_fwd", g2.data_ptr(), act.data_ptr(), g2.shape[0], I, N.stream_ptr())
        ctx.save_for_backward(g2)
        ctx.shape = gu.shape
        return act.view(*gu.shape[:-1], I)

    @staticmethod
    def backward(ctx, dact):
        (g2,) = ctx.saved_tensors
        I = g2.shape[1] // 2
        d2 = dact.reshape(-1, I)
        if not d2.is_contiguous():
            d2 = d2.contiguous()
        dgu = torch.empty_like(g2)
        N.call("stb_swiglu_bwd", d2.data_ptr(), g2.data_ptr(), dgu.data_ptr(), g2.shape[0], I, N.stream_ptr())
        return dgu.view(ctx.shape)


def swiglu(gu: torch.Tensor) -> torch.Tensor:
    """``silu(gate) * up`` on the fused gate|up GEMM output: the sm_100a kernel for contiguous bf16 CUDA tensors, the oracle otherwise."""
    if N.use_native(gu) and gu.dtype == torch.bfloat16 and gu.is_contiguous() and (gu.shape[-1] // 2) % 8 == 0:
        return _SwiGLU.apply(gu)
    return swiglu_ref(gu)
