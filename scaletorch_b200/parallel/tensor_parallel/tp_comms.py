"""Autograd-aware TP collectives (reference: ``scaletorch/parallel/tensor_parallel/tp_comms.py:1-360``).

These are the c10d (NCCL / gloo) implementations -- the measured baseline and the CPU path.  The
fused all-gather→GEMM / GEMM→reduce-scatter kernels live in ``scaletorch_b200.parallel.fused`` and
are selected by ``ParallelContext.fused_comm`` inside the TP layers.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def _ws(group) -> int:
    return 1 if group is None else dist.get_world_size(group)


class _CopyToRegion(torch.autograd.Function):
    """Identity forward, all-reduce backward (Megatron "f")."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        return x

    @staticmethod
    def backward(ctx, g):
        if _ws(ctx.group) > 1:
            g = g.contiguous()
            dist.all_reduce(g, group=ctx.group)
        return g, None


class _ReduceFromRegion(torch.autograd.Function):
    """All-reduce forward, identity backward (Megatron "g").  No defensive clone: the input is a
    fresh GEMM output owned by the caller (the reference clones, ``tp_comms.py:144``)."""

    @staticmethod
    def forward(ctx, x, group):
        if _ws(group) > 1:
            x = x.contiguous()
            dist.all_reduce(x, group=group)
        return x

    @staticmethod
    def backward(ctx, g):
        return g, None


class _GatherLastDim(torch.autograd.Function):
    """All-gather along the last dim forward, take the local slice backward."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        ws = _ws(group)
        if ws == 1:
            return x
        n = x.shape[-1]
        x2 = x.contiguous().view(-1, n)
        out = torch.empty((ws * x2.shape[0], n), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(out, x2, group=group)
        out = out.view(ws, x2.shape[0], n).permute(1, 0, 2).reshape(x2.shape[0], ws * n)
        return out.view(*x.shape[:-1], ws * n)

    @staticmethod
    def backward(ctx, g):
        ws = _ws(ctx.group)
        if ws == 1:
            return g, None
        r = dist.get_rank(ctx.group)
        n = g.shape[-1] // ws
        return g[..., r * n:(r + 1) * n].contiguous(), None


class _AllReduceSum(torch.autograd.Function):
    """Sum across the group in forward *and* backward-identity (each rank's input gets the upstream
    gradient); used by the vocab-parallel loss."""

    @staticmethod
    def forward(ctx, x, group):
        x = x.contiguous().clone()
        if _ws(group) > 1:
            dist.all_reduce(x, group=group)
        return x

    @staticmethod
    def backward(ctx, g):
        return g, None


class _AllGatherDim0(torch.autograd.Function):
    """Sequence-parallel entry: all-gather along dim 0 forward, reduce-scatter backward
    (reference ``sequence_parallel/sp_comms.py:31-61``, there along dim 1 with a ``torch.cat``)."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        ws = _ws(group)
        if ws == 1:
            return x
        x = x.contiguous()
        out = torch.empty((ws * x.shape[0],) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(out, x, group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        ws = _ws(ctx.group)
        if ws == 1:
            return g, None
        return _reduce_scatter_dim0(g.contiguous(), ctx.group), None


def _reduce_scatter_dim0(x: torch.Tensor, group) -> torch.Tensor:
    ws = _ws(group)
    out = torch.empty((x.shape[0] // ws,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
    if dist.get_backend(group) == "gloo":  # gloo has no reduce_scatter
        dist.all_reduce(x, group=group)
        r = dist.get_rank(group)
        out.copy_(x[r * out.shape[0]:(r + 1) * out.shape[0]])
    else:
        dist.reduce_scatter_tensor(out, x, group=group)
    return out


class _ReduceScatterDim0(torch.autograd.Function):
    """Sequence-parallel exit: reduce-scatter along dim 0 forward, all-gather backward
    (reference ``sp_comms.py:64-94``)."""

    @staticmethod
    def forward(ctx, x, group):
        ctx.group = group
        if _ws(group) == 1:
            return x
        return _reduce_scatter_dim0(x.contiguous(), group)

    @staticmethod
    def backward(ctx, g):
        ws = _ws(ctx.group)
        if ws == 1:
            return g, None
        g = g.contiguous()
        out = torch.empty((ws * g.shape[0],) + tuple(g.shape[1:]), dtype=g.dtype, device=g.device)
        dist.all_gather_into_tensor(out, g, group=ctx.group)
        return out, None


def copy_to_tensor_parallel_region(x, group):
    """Identity in forward, all-reduce of the gradient in backward (input of a column-parallel layer)."""
    return _CopyToRegion.apply(x, group)


def reduce_from_tensor_parallel_region(x, group):
    """All-reduce in forward, identity in backward (output of a row-parallel layer)."""
    return _ReduceFromRegion.apply(x, group)


def gather_from_tensor_parallel_region(x, group):
    """All-gather along the last dimension in forward, this rank's slice of the gradient in backward."""
    return _GatherLastDim.apply(x, group)


def all_reduce_sum_autograd(x, group):
    """Sum over the group in forward; the backward hands every rank the upstream gradient unchanged (vocabulary-parallel loss statistics)."""
    return _AllReduceSum.apply(x, group)


def gather_from_sequence_parallel_region(x, group):
    """Sequence-parallel entry: all-gather along dim 0 in forward, reduce-scatter in backward."""
    return _AllGatherDim0.apply(x, group)


def reduce_scatter_to_sequence_parallel_region(x, group):
    """Sequence-parallel exit: reduce-scatter along dim 0 in forward, all-gather in backward."""
    return _ReduceScatterDim0.apply(x, group)


def split_along_dim0(x: torch.Tensor, group) -> torch.Tensor:
    """Local token shard of a replicated ``[T, ...]`` tensor (no communication)."""
    ws = _ws(group)
    if ws == 1:
        return x
    r = dist.get_rank(group)
    n = x.shape[0] // ws
    return x[r * n:(r + 1) * n]
