"""Grouped (ragged) expert GEMMs for MoE layers.

``x_sorted [R, K]`` holds token rows grouped by local expert (``tokens_per_expert[e]`` consecutive rows
for expert ``e``); ``weight [E, N, K]`` is the stacked expert weight.  Computes
``y[rows of e] = x[rows of e] · weight[e]ᵀ`` with the matching dgrad / wgrad.

Reference: the per-expert Python loop with ``nonzero`` → gather → 3 ``nn.Linear`` → ``scatter_`` in
``scaletorch/models/model_qwen3_moe.py:129-171`` (and the Ascend ``npu_grouped_matmul`` wrapper the
authors wanted, ``scaletorch/models/npu_patch.py:94-127``).

Paths: (a) the device-scheduled tcgen05 grouped GEMM (``stb_grouped_gemm_bf16``: one launch, expert
offsets read on the device, no host sync) when the library provides it; (b) otherwise one tcgen05 GEMM
per expert on CUDA / a PyTorch matmul per expert on CPU, driven by host-side offsets.
"""
from __future__ import annotations

import torch

from .. import _native as N
from .activation import swiglu
from .gemm import gemm


def expert_counts(expert_ids: torch.Tensor, num_experts: int) -> torch.Tensor:
    """``bincount(expert_ids, minlength=num_experts)`` without the host synchronisation ``torch.bincount`` performs on
    CUDA to size its output (it reads ``max(input)`` back): the output size is static here."""
    flat = expert_ids.reshape(-1)
    out = torch.zeros(num_experts, dtype=torch.int64, device=flat.device)
    return out.scatter_add_(0, flat, torch.ones_like(flat, dtype=torch.int64))


def _offsets(tokens_per_expert: torch.Tensor) -> list[int]:
    c = tokens_per_expert.tolist()  # host sync (path b only)
    off = [0]
    for n in c:
        off.append(off[-1] + int(n))
    return off


def native_grouped_available() -> bool:
    """True when the in-tree library exports the grouped (per-expert) GEMM."""
    lib = N.load()
    return lib is not None and hasattr(lib, "stb_grouped_gemm_bf16")


class _GroupedLinear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, tokens_per_expert):
        E, Nn, K = weight.shape
        off = _offsets(tokens_per_expert)
        y = torch.empty((x.shape[0], Nn), dtype=x.dtype, device=x.device)
        for e in range(E):
            a, b = off[e], off[e + 1]
            if b > a:
                gemm(x[a:b], weight[e], out=y[a:b])
        if off[-1] < x.shape[0]:
            y[off[-1]:].zero_()
        ctx.save_for_backward(x, weight)
        ctx.off = off
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        off = ctx.off
        E, Nn, K = weight.shape
        dy = dy.contiguous()
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        dw = None
        mg = getattr(weight, "main_grad", None)
        if ctx.needs_input_grad[1]:
            if mg is not None:
                dwv = mg.view(E, Nn, K)
                acc = bool(getattr(weight, "grad_written", False))
            else:
                dwv = torch.zeros((E, Nn, K), dtype=torch.float32 if x.is_cuda else weight.dtype, device=x.device)
                acc = False
        for e in range(E):
            a, b = off[e], off[e + 1]
            if b > a:
                if dx is not None:
                    gemm(dy[a:b], weight[e], b_mn=True, out=dx[a:b])
                if ctx.needs_input_grad[1]:
                    gemm(dy[a:b], x[a:b], a_mn=True, b_mn=True, out=dwv[e], accumulate=acc)
            elif ctx.needs_input_grad[1] and not acc:
                dwv[e].zero_()
        if dx is not None and off[-1] < x.shape[0]:
            dx[off[-1]:].zero_()
        if ctx.needs_input_grad[1]:
            if mg is not None:
                weight.grad_written = True
                hook = getattr(weight, "grad_ready_hook", None)
                if hook is not None:
                    hook()
            else:
                dw = dwv.to(weight.dtype)
        return dx, dw, None


def _use_native(x_sorted, *weights) -> bool:
    if not (native_grouped_available() and N.use_native(x_sorted) and x_sorted.dtype == torch.bfloat16):
        return False
    from .grouped_native import supported

    return all(supported(w) for w in weights)


def grouped_linear(x_sorted: torch.Tensor, weight: torch.Tensor, tokens_per_expert: torch.Tensor) -> torch.Tensor:
    """``y[rows of expert e] = x[rows of expert e] @ weight[e].T`` for rows sorted by expert (``tokens_per_expert`` gives the segment lengths): the tcgen05 grouped GEMM on CUDA, a per-expert loop as the oracle."""
    if _use_native(x_sorted, weight):
        from .grouped_native import grouped_linear_native

        return grouped_linear_native(x_sorted, weight, tokens_per_expert)
    return _GroupedLinear.apply(x_sorted, weight, tokens_per_expert)


def grouped_mlp(x_sorted, tokens_per_expert, gate_up_proj, down_proj) -> torch.Tensor:
    """SwiGLU expert MLP over expert-sorted rows: ``down(silu(gate(x)) * up(x))`` per expert."""
    if _use_native(x_sorted, gate_up_proj, down_proj):
        from .grouped_native import grouped_mlp_native

        return grouped_mlp_native(x_sorted, tokens_per_expert, gate_up_proj, down_proj)
    gu = grouped_linear(x_sorted, gate_up_proj, tokens_per_expert)
    return grouped_linear(swiglu(gu), down_proj, tokens_per_expert)
