"""Expert-parallel token dispatch / combine.

Reference: ``scaletorch/parallel/expert_parallel/ep_comms.py:30-171`` (``all_to_all``,
``dispatch_tokens``, ``gather_tokens``).  Two things are fixed relative to the reference:

* the all-to-all is an autograd function, so gradients flow from expert outputs back to the hidden
  states and to the router (the reference uses raw ``dist.all_to_all_single`` → no gradient through
  the MoE branch);
* routing metadata (expert ids, gate weights) is not shipped as separate all-to-alls: only the
  per-(rank, expert) counts are exchanged, ids are reconstructed from counts on the receiver and gate
  weights are applied on the *sender* after the return trip (3 collectives instead of 5 per layer).

This module is the c10d (NCCL / gloo) path and the numerical baseline; the fused
dispatch→grouped-GEMM→combine kernels over symmetric memory replace it on B200 when available
(``scaletorch_b200.parallel.fused``).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ...ops.grouped import expert_counts


def _a2a_single(out: torch.Tensor, inp: torch.Tensor, out_splits, in_splits, group) -> None:
    if dist.get_backend(group) == "gloo":
        # gloo lacks all_to_all_single with uneven splits on some builds: emulate with p2p
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        in_off = [0]
        for s in in_splits:
            in_off.append(in_off[-1] + s)
        out_off = [0]
        for s in out_splits:
            out_off.append(out_off[-1] + s)
        reqs = []
        for peer in range(world):
            src = inp[in_off[peer]:in_off[peer + 1]]
            dst = out[out_off[peer]:out_off[peer + 1]]
            if peer == rank:
                dst.copy_(src)
                continue
            g = dist.get_global_rank(group, peer)
            if src.numel() > 0:
                reqs.append(dist.isend(src.contiguous(), g, group=group))
            if dst.numel() > 0:
                reqs.append(dist.irecv(dst, g, group=group))
        for r in reqs:
            r.wait()
    else:
        dist.all_to_all_single(out, inp, out_splits, in_splits, group=group)


class _AllToAll(torch.autograd.Function):
    """Variable-split all-to-all along dim 0 with the mirrored all-to-all as backward."""

    @staticmethod
    def forward(ctx, x, out_splits, in_splits, group):
        ctx.group, ctx.out_splits, ctx.in_splits = group, out_splits, in_splits
        x = x.contiguous()
        out = torch.empty((sum(out_splits),) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        _a2a_single(out, x, out_splits, in_splits, group)
        return out

    @staticmethod
    def backward(ctx, g):
        g = g.contiguous()
        out = torch.empty((sum(ctx.in_splits),) + tuple(g.shape[1:]), dtype=g.dtype, device=g.device)
        _a2a_single(out, g, ctx.in_splits, ctx.out_splits, ctx.group)
        return out, None, None, None


def all_to_all(x: torch.Tensor, output_split_sizes, input_split_sizes, group) -> torch.Tensor:
    """Differentiable ``all_to_all_single`` (rows of ``x`` are split by ``input_split_sizes``)."""
    if group is None or dist.get_world_size(group) == 1:
        return x
    return _AllToAll.apply(x, list(output_split_sizes), list(input_split_sizes), group)


def exchange_counts(counts: torch.Tensor, group) -> torch.Tensor:
    """counts ``[ep, E_local]`` (rows this rank sends to each (rank, local expert)) →
    ``[ep, E_local]`` rows this rank *receives* from each source rank for each of its experts."""
    out = torch.empty_like(counts)
    ws = dist.get_world_size(group)
    n = counts.shape[1]
    _a2a_single(out.view(-1), counts.contiguous().view(-1), [n] * ws, [n] * ws, group)
    return out


def dispatch_tokens(hidden_states: torch.Tensor, topk_indices: torch.Tensor, num_experts: int, ep_size: int,
                    group):
    """Routes every ``(token, k)`` pair to the EP rank that owns its expert.

    Returns ``(recv_tokens [R, H] sorted by local expert, tokens_per_local_expert [E_local], state)``
    where ``state`` carries what :func:`gather_tokens` needs for the return trip.
    """
    T, H = hidden_states.shape
    k = topk_indices.shape[1]
    e_local = num_experts // ep_size
    flat = topk_indices.reshape(-1)
    sort_idx = torch.argsort(flat, stable=True)          # (token,k) pairs ordered by global expert id
    token_of = torch.div(sort_idx, k, rounding_mode="floor")
    send = hidden_states.index_select(0, token_of)       # [T*k, H], differentiable gather
    counts = expert_counts(flat, num_experts).view(ep_size, e_local)
    recv_counts = exchange_counts(counts, group)          # [src_rank, local expert]
    send_splits = counts.sum(1).tolist()                  # host sync (c10d needs host-side splits)
    recv_splits = recv_counts.sum(1).tolist()
    recv = all_to_all(send, recv_splits, send_splits, group)   # ordered by (src rank, local expert)
    # re-order (src, expert)-major → expert-major so each expert's rows are contiguous
    R = recv.shape[0]
    rc = recv_counts.reshape(-1)                          # [ep*e_local] in (src, expert) order
    seg_expert = torch.arange(e_local, device=recv.device).repeat(ep_size)
    row_expert = torch.repeat_interleave(seg_expert, rc, output_size=R)
    local_sort = torch.argsort(row_expert, stable=True)
    recv_sorted = recv.index_select(0, local_sort)
    tokens_per_expert = recv_counts.sum(0)
    state = dict(sort_idx=sort_idx, token_of=token_of, local_sort=local_sort, send_splits=send_splits,
                 recv_splits=recv_splits, num_tokens=T, top_k=k)
    return recv_sorted, tokens_per_expert, state


def gather_tokens(expert_output: torch.Tensor, topk_weights: torch.Tensor, state: dict, group) -> torch.Tensor:
    """Return trip of :func:`dispatch_tokens`: un-sort, all-to-all back, weight by the gate and sum the
    ``k`` contributions of each token → ``[T, H]``."""
    inv = torch.empty_like(state["local_sort"])
    inv[state["local_sort"]] = torch.arange(inv.numel(), device=inv.device)
    back = expert_output.index_select(0, inv)             # (src, expert)-major again
    ret = all_to_all(back, state["send_splits"], state["recv_splits"], group)  # [T*k, H] in sort order
    w = topk_weights.reshape(-1).index_select(0, state["sort_idx"]).to(ret.dtype)
    ret = ret * w.unsqueeze(-1)
    out = torch.zeros((state["num_tokens"], ret.shape[1]), dtype=ret.dtype, device=ret.device)
    return out.index_add(0, state["token_of"], ret)
