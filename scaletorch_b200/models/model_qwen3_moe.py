"""Qwen3 Mixture-of-Experts with expert parallelism (reference:
``scaletorch/models/model_qwen3_moe.py:30-409``).

Qwen3-30B-A3B: 48 layers, hidden 2048, 32 heads / 4 KV heads, head_dim 128, 128 experts, top-8,
``moe_intermediate_size`` 768, ``norm_topk_prob``.

B200-first differences:
* experts are stacked tensors (``gate_up_proj [E_local, 2I, H]``, ``down_proj [E_local, H, I]``)
  consumed by a ragged grouped GEMM over expert-sorted rows -- no ``nn.ModuleList`` / Python loop with
  ``nonzero`` per expert (reference ``:129-171``);
* dispatch/combine are differentiable (the reference's raw all-to-all drops the gradient, §2.7 #7)
  and the load-balancing auxiliary loss is exposed through ``get_aux_loss()`` *and* added by the
  trainer (the reference never adds it);
* EP ranks consume distinct data by default (``ProcessGroupManager.ep_replicate_data``).
"""
from __future__ import annotations

from typing import Any

import torch
import torch.nn as nn

from .. import ops
from ..ops.grouped import expert_counts, grouped_mlp
from ..parallel.context import ParallelContext
from ..parallel.expert_parallel.ep_comms import dispatch_tokens, gather_tokens
from .transformer import MLP, DecoderLM


class MoERouter(nn.Module):
    """Top-k gating: gate GEMM → fp32 logits → top-k → softmax over the k → (renorm) → Switch aux loss."""

    def __init__(self, hidden_size: int, num_experts: int, top_k: int, norm_topk_prob: bool = True,
                 aux_loss_coef: float = 0.001):
        super().__init__()
        self.num_experts, self.top_k = num_experts, top_k
        self.norm_topk_prob, self.aux_loss_coef = norm_topk_prob, aux_loss_coef
        self.gate = nn.Linear(hidden_size, num_experts, bias=False)

    def reset_parameters(self) -> None:
        if self.gate.weight.device.type != "meta":
            nn.init.normal_(self.gate.weight, std=0.02)

    def forward(self, x: torch.Tensor):
        """x ``[T, H]`` → ``(weights [T,k] fp32, indices [T,k], aux_loss scalar)``.

        Gate GEMM, then ONE routing kernel (``ops.moe_routing``): top-k, softmax over the k selected logits (which already
        sums to one, so ``norm_topk_prob`` changes nothing), per-expert counts, the position of every pair inside its
        expert and the Switch aux loss.  Counts and positions are kept in ``self.plan`` for the expert-parallel dispatch."""
        from ..ops.moe_routing import route

        logits = ops.linear(x, self.gate.weight)
        weights, idx, pos, counts, aux = route(logits, self.top_k, self.aux_loss_coef, self.training)
        self.plan = (pos, counts)
        return weights, idx, aux


class MoEExperts(nn.Module):
    """``E_local`` SwiGLU experts as stacked weights for the grouped GEMM."""

    def __init__(self, num_local_experts: int, hidden_size: int, intermediate_size: int):
        super().__init__()
        self.num_local_experts = num_local_experts
        self.hidden_size, self.intermediate_size = hidden_size, intermediate_size
        self.gate_up_proj = nn.Parameter(torch.empty(num_local_experts, 2 * intermediate_size, hidden_size))
        self.down_proj = nn.Parameter(torch.empty(num_local_experts, hidden_size, intermediate_size))
        self.reset_parameters()

    def reset_parameters(self) -> None:
        if self.gate_up_proj.device.type != "meta":
            nn.init.normal_(self.gate_up_proj, std=0.02)
            nn.init.normal_(self.down_proj, std=0.02)

    def forward(self, x_sorted: torch.Tensor, tokens_per_expert: torch.Tensor) -> torch.Tensor:
        """x_sorted ``[R, H]`` with rows grouped by local expert; ``tokens_per_expert [E_local]`` (device)."""
        return grouped_mlp(x_sorted, tokens_per_expert, self.gate_up_proj, self.down_proj)


class MoELayer(nn.Module):
    """Sparse MLP block: router -> top-k experts -> weighted sum.  Local experts run as a sort + grouped GEMM; with expert parallelism tokens travel through the peer-memory dispatch / combine kernels (or the differentiable all-to-all path on c10d)."""
    def __init__(self, config: Any, ctx: ParallelContext):
        super().__init__()
        self.ctx = ctx
        self.num_experts = config.num_experts
        self.top_k = config.num_experts_per_tok
        self.hidden_size = config.hidden_size
        self.moe_intermediate_size = config.moe_intermediate_size
        self.ep_size, self.ep_rank = ctx.ep_size, ctx.ep_rank
        if self.num_experts % self.ep_size != 0:
            raise ValueError(f"num_experts ({self.num_experts}) must be divisible by ep size ({self.ep_size})")
        self.experts_per_rank = self.num_experts // self.ep_size
        self.router = MoERouter(config.hidden_size, self.num_experts, self.top_k,
                                norm_topk_prob=getattr(config, "norm_topk_prob", True),
                                aux_loss_coef=getattr(config, "router_aux_loss_coef", 0.001))
        self.experts = MoEExperts(self.experts_per_rank, config.hidden_size, config.moe_intermediate_size)
        self._aux_loss: torch.Tensor | None = None
        self.use_fused_ep = True                     # peer-memory dispatch/combine when the fabric allows it
        self.ep_capacity_factor = getattr(config, "ep_capacity_factor", None)
        self.dropped_pairs: torch.Tensor | None = None   # device counter of over-capacity pairs (fused EP path)

    def reset_parameters(self) -> None:
        self.router.reset_parameters()
        self.experts.reset_parameters()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x ``[T, H]`` → ``[T, H]``.  Under TP the input is replicated (or gathered from the SP shard by
        the caller) and every TP rank computes the same MoE output -- experts are not TP-sharded, as in
        the reference (``tensor_parallel.py:132-138``)."""
        ctx = self.ctx
        if ctx.sp:
            # every TP rank runs the whole MoE branch on all tokens (experts are not TP-sharded) and keeps its token shard
            # of the result; because each rank therefore sees the *complete* output gradient, the gradient w.r.t. the
            # gathered input is complete on every rank too: the backward of this gather is a slice, not the
            # reduce-scatter of the Megatron-SP gather (which would count the branch tp times)
            x = _GatherKeepGrad.apply(x, ctx.tp_group)
        weights, idx, aux = self.router(x)
        self._aux_loss = aux
        if self.ep_size > 1 and self.use_fused_ep and self._fused_ep_ok(x):
            from ..parallel.expert_parallel import fused_ep

            out, st = fused_ep.fused_moe_ep(x, weights, idx, self.experts.gate_up_proj, self.experts.down_proj,
                                            self.num_experts, ctx.ep_group, self.ep_capacity_factor,
                                            plan=self.router.plan)
            self.dropped_pairs = st.dropped if self.dropped_pairs is None else self.dropped_pairs + st.dropped
        elif self.ep_size > 1:
            idx = idx.long()
            recv, tokens_per_expert, state = dispatch_tokens(x, idx, self.num_experts, self.ep_size, ctx.ep_group)
            y = self.experts(recv, tokens_per_expert)
            out = gather_tokens(y, weights, state, ctx.ep_group)
        else:
            out = self._forward_local(x, weights, idx)
        out = out.to(x.dtype)
        if ctx.sp:
            # every TP rank holds the full, identical output: keep the local token shard
            out = _SplitKeepGrad.apply(out, ctx.tp_group)
        return out

    def _fused_ep_ok(self, x) -> bool:
        if not x.is_cuda:
            return False
        from ..parallel.expert_parallel import fused_ep

        return fused_ep.available(self.ctx.ep_group, x, self.experts.gate_up_proj, self.experts.down_proj)

    def _forward_local(self, x, weights, idx):
        T, H = x.shape
        k = idx.shape[1]
        flat = idx.reshape(-1).long()
        sort_idx = torch.argsort(flat, stable=True)
        token_of = torch.div(sort_idx, k, rounding_mode="floor")
        xs = x.index_select(0, token_of)
        counts = expert_counts(flat, self.num_experts)
        ys = self.experts(xs, counts)
        w = weights.reshape(-1).index_select(0, sort_idx).to(ys.dtype)
        ys = ys * w.unsqueeze(-1)
        return torch.zeros((T, H), dtype=ys.dtype, device=ys.device).index_add(0, token_of, ys)


class _GatherKeepGrad(torch.autograd.Function):
    """All-gather the TP ranks' token shards along dim 0; the backward keeps this rank's slice of the (already
    complete) gradient.  Counterpart of :class:`_SplitKeepGrad` around the TP-replicated MoE branch."""

    @staticmethod
    def forward(ctx, x, group):
        import torch.distributed as dist

        ctx.group = group
        ws = dist.get_world_size(group)
        x = x.contiguous()
        out = torch.empty((ws * x.shape[0],) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        dist.all_gather_into_tensor(out, x, group=group)
        return out

    @staticmethod
    def backward(ctx, g):
        import torch.distributed as dist

        ws, r = dist.get_world_size(ctx.group), dist.get_rank(ctx.group)
        n = g.shape[0] // ws
        return g[r * n:(r + 1) * n].contiguous(), None


class _SplitKeepGrad(torch.autograd.Function):
    """Take this TP rank's token shard of a tensor that is replicated across the TP group; the backward
    all-gathers the shard gradients (each rank contributes its own rows)."""

    @staticmethod
    def forward(ctx, x, group):
        import torch.distributed as dist

        ctx.group = group
        ws, r = dist.get_world_size(group), dist.get_rank(group)
        n = x.shape[0] // ws
        return x[r * n:(r + 1) * n].contiguous()

    @staticmethod
    def backward(ctx, g):
        import torch.distributed as dist

        ws = dist.get_world_size(ctx.group)
        out = torch.empty((ws * g.shape[0],) + tuple(g.shape[1:]), dtype=g.dtype, device=g.device)
        dist.all_gather_into_tensor(out, g.contiguous(), group=ctx.group)
        return out, None


class Qwen3MoE(DecoderLM):
    """Qwen3-MoE decoder: Qwen3 attention with a ``MoELayer`` as the feed-forward of every sparse layer; ``get_aux_loss()`` sums the routers' load-balancing losses."""
    model_type = "qwen3_moe"
    qk_norm = True
    default_rope_theta = 1000000.0
    attn_init_std = 0.02

    def _head_dim(self, config) -> int:
        return getattr(config, "head_dim", None) or config.hidden_size // config.num_attention_heads

    def _is_moe_layer(self, config, layer_idx: int) -> bool:
        step = getattr(config, "decoder_sparse_step", 1) or 1
        mlp_only = getattr(config, "mlp_only_layers", None) or []
        return getattr(config, "num_experts", 0) > 0 and layer_idx not in mlp_only and (layer_idx + 1) % step == 0

    def _make_ffn(self, config, layer_idx: int):
        if self._is_moe_layer(config, layer_idx):
            return MoELayer(config, self.ctx)
        return None

    def get_aux_loss(self) -> torch.Tensor | None:
        """Sum of the router load-balancing losses of the last forward (None if there was none)."""
        total = None
        for layer in self.decoder_layers:
            moe = getattr(layer, "moe", None)
            if moe is not None and moe._aux_loss is not None:
                total = moe._aux_loss if total is None else total + moe._aux_loss
        return total


__all__ = ["MoERouter", "MoEExperts", "MoELayer", "Qwen3MoE", "MLP"]
