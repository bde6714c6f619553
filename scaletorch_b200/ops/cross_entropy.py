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
    return F.cross_entropy(logits.float().view(-1, logits.shape[-1]), targets.reshape(-1), ignore_index=ignore_index)


class _LinearCrossEntropy(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, targets, loss_scale, ignore_index, tp_group, vocab_start):
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        T = x2.shape[0]
        V = weight.shape[0]
        tgt = targets.reshape(-1).contiguous()
        logits = gemm(x2, weight)  # [T, V_local] bf16
        stats = torch.empty((T, 3), dtype=torch.float32, device=x.device)
        N.call("stb_ce_stats", logits.data_ptr(), tgt.data_ptr(), stats.data_ptr(), T, V, logits.stride(0),
               vocab_start, N.stream_ptr())
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
        N.call("stb_ce_grad", logits.data_ptr(), tgt.data_ptr(), lse.data_ptr(), T, V, logits.stride(0),
               vocab_start, 1.0, scale.data_ptr(), ignore_index, N.stream_ptr())
        dx = gemm(logits, weight, b_mn=True).view(xshape) if ctx.needs_input_grad[0] else None
        dw = wgrad_into(weight, logits, x2) if ctx.needs_input_grad[1] else None
        return dx, dw, None, None, None, None, None


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
