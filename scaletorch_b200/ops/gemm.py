"""bf16 GEMM on tcgen05 tensor cores + the autograd ``linear`` built on it.

``gemm`` computes ``C[M,N] (+)= A · Bᵀ`` where each operand is either K-major
(row-major ``[rows, K]``) or MN-major (row-major ``[K, rows]``), so the three
GEMMs of a Linear layer (fprop / dgrad / wgrad) all run without transposed
copies.  Reference behaviour: ``F.linear`` in
``scaletorch/parallel/tensor_parallel/tensor_parallel.py:240,352`` and the
hand-written linear backward in ``tp_comms.py:265-320``.

Weight gradients follow the *main-grad* convention of the framework's flat
gradient store (see ``scaletorch_b200.runtime.flat_params``): when a parameter
carries ``main_grad`` (an fp32 view into the flat gradient buffer) the wgrad
GEMM writes/accumulates straight into it from its epilogue and autograd sees
``None`` for that parameter.
"""
from __future__ import annotations

import torch

from .. import _native as N


def gemm(
    a: torch.Tensor,
    b: torch.Tensor,
    *,
    a_mn: bool = False,
    b_mn: bool = False,
    out: torch.Tensor | None = None,
    out_dtype: torch.dtype | None = None,
    accumulate: bool = False,
    force_bn: int = 0,
) -> torch.Tensor:
    """``out[M,N] (+)= A · Bᵀ``.

    a: ``[M,K]`` (``a_mn=False``) or ``[K,M]`` (``a_mn=True``); b: ``[N,K]`` or ``[K,N]``.
    """
    assert a.dim() == 2 and b.dim() == 2
    M, K = (a.shape[1], a.shape[0]) if a_mn else a.shape
    Nn, Kb = (b.shape[1], b.shape[0]) if b_mn else b.shape
    assert K == Kb, f"K mismatch {K} vs {Kb}"
    if out is None:
        assert not accumulate
        out = torch.empty((M, Nn), dtype=out_dtype or a.dtype, device=a.device)
    if not N.use_native(a):
        af = (a.t() if a_mn else a).float()
        bf = (b.t() if b_mn else b).float()
        res = af @ bf.t()
        if accumulate:
            out.add_(res.to(out.dtype))
        else:
            out.copy_(res.to(out.dtype))
        return out
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert a.stride(1) == 1 and b.stride(1) == 1 and out.stride(1) == 1
    N.call(
        "stb_gemm_bf16", a.data_ptr(), b.data_ptr(), out.data_ptr(), M, Nn, K,
        a.stride(0), b.stride(0), out.stride(0), int(a_mn), int(b_mn),
        int(out.dtype == torch.float32), int(accumulate), force_bn, N.stream_ptr(),
    )
    return out


def gemm_pair(a, b, out, *, a_mn: bool = False, b_mn: bool = False, accumulate: bool = False, bn: int = 0,
              splits: int = 0):
    """The CTA-pair (``cta_group::2``) kernel called directly, with an explicit tile width (128 / 256) and k-split (0 = let the library choose).
    ``gemm`` reaches the same kernel through ``stb_gemm_bf16``'s own dispatch; this entry exists for tests and tuning."""
    M, K = (a.shape[1], a.shape[0]) if a_mn else a.shape
    Nn = b.shape[1] if b_mn else b.shape[0]
    N.call("stb_gemm_bf16_2cta", a.data_ptr(), b.data_ptr(), out.data_ptr(), M, Nn, K, a.stride(0), b.stride(0),
           out.stride(0), int(a_mn), int(b_mn), int(out.dtype == torch.float32), int(accumulate), bn, splits, N.stream_ptr())
    return out


def wgrad_into(param: torch.Tensor, dy2d: torch.Tensor, x2d: torch.Tensor) -> torch.Tensor | None:
    """dW = dyᵀ·x.  Writes into ``param.main_grad`` when present (returns None), else returns dW."""
    mg = getattr(param, "main_grad", None)
    if mg is not None:
        acc = bool(getattr(param, "grad_written", False))
        gemm(dy2d, x2d, a_mn=True, b_mn=True, out=mg.view(param.shape), accumulate=acc)
        param.grad_written = True
        hook = getattr(param, "grad_ready_hook", None)
        if hook is not None:
            hook()
        return None
    return gemm(dy2d, x2d, a_mn=True, b_mn=True, out_dtype=param.dtype)


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight):
        ctx.save_for_backward(x, weight)
        x2d = x.reshape(-1, x.shape[-1])
        y = gemm(x2d, weight)
        return y.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        dy2d = dy.reshape(-1, dy.shape[-1])
        if not dy2d.is_contiguous():
            dy2d = dy2d.contiguous()
        x2d = x.reshape(-1, x.shape[-1])
        dx = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dy2d, weight, b_mn=True).view_as(x)
        dw = wgrad_into(weight, dy2d, x2d) if ctx.needs_input_grad[1] else None
        return dx, dw


def linear(x: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor | None = None) -> torch.Tensor:
    """``y = x Wᵀ (+ b)``; tcgen05 GEMMs on CUDA bf16, ``F.linear`` otherwise."""
    if N.use_native(x) and x.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16:
        y = _Linear.apply(x, weight)
        return y if bias is None else y + bias
    return torch.nn.functional.linear(x, weight, bias)
