"""MoE routing as device kernels (``csrc/moe_router.cu``): gate logits → top-k ids, gate weights, per-expert counts, the
position of every (token, k) pair inside its expert's segment and the load-balancing loss -- one kernel forward, one
backward, no library sort and no host synchronisation.

Reference: ``torch.topk`` / ``F.softmax`` / bincount + eager aux loss in ``scaletorch/models/model_qwen3_moe.py:30-75``.
The PyTorch implementation below is the CPU path and the numerical oracle of the kernels.
"""
from __future__ import annotations

import ctypes

import torch
import torch.nn.functional as F

from .. import _native as N

_P, _I, _L, _F = ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_float
N.register("stb_moe_router_fwd", [_P, _L, _I, _I, _P, _P, _P, _P, _P, _P])
N.register("stb_moe_router_bwd", [_P, _P, _P, _P, _P, _P, _F, _L, _I, _I, _P, _P])


def native_router_available(logits: torch.Tensor, k: int) -> bool:
    """True when the fused router kernel applies: bf16 CUDA logits, at most 256 experts, ``k <= 16``."""
    lib = N.load() if logits.is_cuda else None
    return (lib is not None and hasattr(lib, "stb_moe_router_fwd") and N.use_native(logits)
            and logits.dtype == torch.bfloat16 and logits.shape[1] <= 256 and k <= 16)


def route_ref(logits: torch.Tensor, k: int, aux_coef: float, training: bool):
    """Oracle: ``(weights [T,k] fp32, idx [T,k] int64, pos [T,k] int64, counts [E] int64, aux scalar)``.
    ``pos`` is the stable position of a pair among the pairs of the same expert (any bijection would do)."""
    T, E = logits.shape
    lf = logits.float()
    topv, idx = torch.topk(lf, k, dim=-1)
    weights = F.softmax(topv, dim=-1)
    flat = idx.reshape(-1)
    counts = torch.zeros(E, dtype=torch.int64, device=logits.device).scatter_add_(0, flat, torch.ones_like(flat))
    order = torch.argsort(flat, stable=True)
    inv = torch.empty_like(order)
    inv[order] = torch.arange(flat.numel(), device=flat.device)
    pos = (inv - (torch.cumsum(counts, 0) - counts)[flat]).view(T, k)
    aux = lf.new_zeros(())
    if training and aux_coef > 0:
        probs = F.softmax(lf, dim=-1)
        aux = aux_coef * E * ((counts.to(probs.dtype) / T) * probs.mean(dim=0)).sum()
    return weights, idx, pos, counts, aux


class _Route(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, k, aux_coef, training):
        T, E = logits.shape
        dev = logits.device
        logits = logits.contiguous()
        weights = torch.empty((T, k), dtype=torch.float32, device=dev)
        idx = torch.empty((T, k), dtype=torch.int32, device=dev)
        pos = torch.empty((T, k), dtype=torch.int32, device=dev)
        acc = torch.zeros(2 * E, dtype=torch.float32, device=dev)          # [counts as int32 bits | psum]
        counts = acc[:E].view(torch.int32)
        psum = acc[E:]
        N.call("stb_moe_router_fwd", logits.data_ptr(), T, E, k, weights.data_ptr(), idx.data_ptr(), pos.data_ptr(),
               counts.data_ptr(), psum.data_ptr(), N.stream_ptr())
        use_aux = bool(training and aux_coef > 0)
        aux = (counts.float() * psum).sum() * (aux_coef * E / (float(T) * float(T))) if use_aux else logits.new_zeros((), dtype=torch.float32)
        ctx.save_for_backward(logits, weights, idx, counts)
        ctx.cfg = (k, aux_coef if use_aux else 0.0)
        ctx.mark_non_differentiable(idx, pos, counts)
        return weights, idx, pos, counts, aux

    @staticmethod
    def backward(ctx, dweights, _di, _dp, _dc, daux):
        logits, weights, idx, counts = ctx.saved_tensors
        k, coef = ctx.cfg
        T, E = logits.shape
        dlogits = torch.empty_like(logits)
        dw = dweights.contiguous().float() if dweights is not None else torch.zeros_like(weights)
        ga = daux.reshape(1).float().contiguous() if (daux is not None and coef > 0) else None
        N.call("stb_moe_router_bwd", logits.data_ptr(), weights.data_ptr(), idx.data_ptr(), dw.data_ptr(),
               counts.data_ptr(), N.ptr(ga), float(coef * E / float(T)), T, E, k, dlogits.data_ptr(), N.stream_ptr())
        return dlogits, None, None, None


def route(logits: torch.Tensor, k: int, aux_coef: float = 0.0, training: bool = True):
    """``logits [T, E]`` → ``(weights [T,k] fp32, idx [T,k], pos [T,k], counts [E], aux)``; int32 ids / positions /
    counts from the kernel, int64 from the oracle."""
    if native_router_available(logits, k):
        return _Route.apply(logits, k, float(aux_coef), bool(training))
    return route_ref(logits, k, aux_coef, training)
