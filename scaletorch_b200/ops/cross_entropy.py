"""Fused LM-head + cross-entropy (optionally vocab-parallel).

The reference materialises full-vocab logits, all-gathers them across TP ranks and runs
``F.cross_entropy`` (``scaletorch/trainer/train_step.py:89-103``,
``scaletorch/parallel/tensor_parallel/tensor_parallel.py:142,248``).  Here the LM-head GEMM
writes bf16 logits once; a statistics kernel produces per-row (max, sum-exp, target logit);
with TP only those three scalars per row are all-reduced (vocab-parallel loss); the backward
overwrites the logits in place with ``(softmax - onehot) * scale`` and feeds them to the
dgrad / wgrad GEMMs.  fp32 logits and the all-gather never exist.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn.functional as F

from .. import _native as N
from .gemm import gemm, wgrad_into


def cross_entropy_ref(logits: torch.Tensor, targets: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """PyTorch oracle: mean cross-entropy of fp32-upcast logits over the targets that are not ``ignore_index``."""
    return F.cross_entropy(logits.float().view(-1, logits.shape[-1]), targets.reshape(-1), ignore_index=ignore_index)


def _chunks(T: int, V: int) -> list[tuple[int, int]]:
    """Column chunks of the vocabulary such that one ``[T, Vc]`` bf16 chunk of logits (plus its weight rows) stays in the
    126 MB L2 between the kernel that writes it and the kernels that read it: forward = GEMM → statistics, backward =
    gradient-in-place → dgrad GEMM → wgrad GEMM.  The logits still exist in HBM once (the backward needs them), but the
    statistics pass and both backward GEMMs read them from L2: ≈3 × T·V·2 bytes of HBM reads saved per step
    (3.7 GB for 4096 tokens of a 151,936-word vocabulary)."""
    import os

    # Off by default: measured on B200 (Qwen3-0.6B, 4096 tokens, V = 151,936; profiles/ce_chunking_r2.md) the 19 chunks
    # cost more in launches, GEMM prologue/epilogue tails and the fp32 dX accumulation than the L2 hits save
    # (27.09 ms/step chunked vs 26.75 ms unchunked).  SCALETORCH_B200_CE_CHUNK_MB=64 turns it on.
    budget = int(os.environ.get("SCALETORCH_B200_CE_CHUNK_MB", "0")) << 20
    if budget <= 0 or T * V * 2 <= 2 * budget:
        return [(0, V)]
    vc = max(1024, (budget // (T * 2)) // 256 * 256)
    return [(c, min(V, c + vc)) for c in range(0, V, vc)]


class _LinearCrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, targets, loss_scale, ignore_index, tp_group, vocab_start):
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        T = x2.shape[0]
        V = weight.shape[0]
        tgt = targets.reshape(-1).contiguous()
        stats = torch.empty((T, 3), dtype=torch.float32, device=x.device)
        chunks = _chunks(T, V) if (V % 8 == 0 and getattr(weight, "main_grad", None) is not None) else [(0, V)]
        if len(chunks) == 1:
            logits = gemm(x2, weight)  # [T, V_local] bf16
            N.call("stb_ce_stats", logits.data_ptr(), tgt.data_ptr(), stats.data_ptr(), T, V, logits.stride(0),
                   vocab_start, N.stream_ptr())
        else:
            logits = torch.empty((T, V), dtype=x2.dtype, device=x.device)
            for i, (c0, c1) in enumerate(chunks):
                lc = logits[:, c0:c1]
                gemm(x2, weight[c0:c1], out=lc)
                N.call("stb_ce_stats_chunk", lc.data_ptr(), tgt.data_ptr(), stats.data_ptr(), T, c1 - c0, V,
                       vocab_start + c0, int(i > 0), N.stream_ptr())
        ctx.chunks = chunks
        if tp_group is not None and dist.get_world_size(tp_group) > 1:
            gmax = stats[:, 0].clone()
            dist.all_reduce(gmax, op=dist.ReduceOp.MAX, group=tp_group)
            st = torch.stack([stats[:, 1] * torch.exp(stats[:, 0] - gmax), stats[:, 2]], dim=1)
            dist.all_reduce(st, group=tp_group)
            stats = torch.stack([gmax, st[:, 0], st[:, 1]], dim=1).contiguous()
        lse = torch.empty(T, dtype=torch.float32, device=x.device)
        loss_rows = torch.empty(T, dtype=torch.float32, device=x.device)
        N.call("stb_ce_finalize", stats.data_ptr(), tgt.data_ptr(), lse.data_ptr(), loss_rows.data_ptr(), T,
               ignore_index, N.stream_ptr())
        n_valid = (tgt != ignore_index).sum().clamp_(min=1).to(torch.float32)
        inv = loss_scale / n_valid
        loss = loss_rows.sum() * inv
        ctx.save_for_backward(x2, weight, tgt, logits, lse, inv)
        ctx.cfg = (ignore_index, vocab_start, x.shape)
        return loss

    @staticmethod
    def backward(ctx, dloss):
        x2, weight, tgt, logits, lse, inv = ctx.saved_tensors
        ignore_index, vocab_start, xshape = ctx.cfg
        T, V = logits.shape
        scale = (dloss.to(torch.float32) * inv).contiguous()
        if len(ctx.chunks) > 1:
            return _LinearCrossEntropy._backward_chunked(ctx, x2, weight, tgt, logits, lse, scale)
        N.call("stb_ce_grad", logits.data_ptr(), tgt.data_ptr(), lse.data_ptr(), T, V, logits.stride(0),
               vocab_start, 1.0, scale.data_ptr(), ignore_index, N.stream_ptr())
        dx = gemm(logits, weight, b_mn=True).view(xshape) if ctx.needs_input_grad[0] else None
        dw = wgrad_into(weight, logits, x2) if ctx.needs_input_grad[1] else None
        return dx, dw, None, None, None, None, None


def _backward_chunked(ctx, x2, weight, tgt, logits, lse, scale):
    """Per column chunk: gradient in place → dgrad (accumulated into an fp32 dX) → wgrad into the chunk's rows of the
    flat fp32 gradient buffer; each chunk is consumed out of L2 right after it is produced."""
    ignore_index, vocab_start, xshape = ctx.cfg
    T, V = logits.shape
    H = x2.shape[1]
    need_dx, need_dw = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
    dx32 = torch.empty((T, H), dtype=torch.float32, device=x2.device) if need_dx else None
    mg = weight.main_grad.view(V, H)
    acc_w = bool(getattr(weight, "grad_written", False))
    for i, (c0, c1) in enumerate(ctx.chunks):
        lc = logits[:, c0:c1]
        N.call("stb_ce_grad", lc.data_ptr(), tgt.data_ptr(), lse.data_ptr(), T, c1 - c0, V, vocab_start + c0, 1.0,
               scale.data_ptr(), ignore_index, N.stream_ptr())
        if need_dx:
            gemm(lc, weight[c0:c1], b_mn=True, out=dx32, accumulate=i > 0)
        if need_dw:
            gemm(lc, x2, a_mn=True, b_mn=True, out=mg[c0:c1], accumulate=acc_w)
    if need_dw:
        weight.grad_written = True
        hook = getattr(weight, "grad_ready_hook", None)
        if hook is not None:
            hook()
    dx = dx32.to(x2.dtype).view(xshape) if need_dx else None
    return dx, None, None, None, None, None, None


_LinearCrossEntropy._backward_chunked = staticmethod(_backward_chunked)


def linear_cross_entropy(
    x: torch.Tensor,
    weight: torch.Tensor,
    targets: torch.Tensor,
    *,
    loss_scale: float = 1.0,
    ignore_index: int = -100,
    tp_group=None,
    vocab_start: int = 0,
) -> torch.Tensor:
    """mean CE of ``x @ weightᵀ`` against ``targets`` times ``loss_scale``.

    ``weight`` may be a vocab shard ``[V/tp, H]`` starting at ``vocab_start`` (then ``tp_group``
    must be given); the returned loss is identical on every TP rank.
    """
    if N.use_native(x) and x.dtype == torch.bfloat16:
        return _LinearCrossEntropy.apply(x, weight, targets, float(loss_scale), int(ignore_index), tp_group, int(vocab_start))
    # oracle / CPU path (also vocab-parallel aware, via autograd-friendly collectives)
    logits = F.linear(x, weight).float()
    tgt = targets.reshape(-1)
    lg = logits.view(-1, logits.shape[-1])
    if tp_group is None or dist.get_world_size(tp_group) == 1:
        return F.cross_entropy(lg, tgt, ignore_index=ignore_index) * loss_scale
    from ..parallel.tensor_parallel.tp_comms import all_reduce_sum_autograd

    V = lg.shape[1]
    gmax = lg.detach().max(dim=1).values
    dist.all_reduce(gmax, op=dist.ReduceOp.MAX, group=tp_group)
    sumexp = all_reduce_sum_autograd(torch.exp(lg - gmax[:, None]).sum(1), tp_group)
    local = tgt - vocab_start
    mask = (local >= 0) & (local < V)
    tl = lg.gather(1, local.clamp(0, V - 1)[:, None]).squeeze(1) * mask
    tl = all_reduce_sum_autograd(tl, tp_group)
    loss_rows = (gmax + torch.log(sumexp) - tl)
    valid = tgt != ignore_index
    return (loss_rows * valid).sum() / valid.sum().clamp(min=1) * loss_scale
