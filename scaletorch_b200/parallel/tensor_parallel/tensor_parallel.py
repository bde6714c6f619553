"""Megatron-style 1-D tensor parallelism with real sequence parallelism.

Reference: ``scaletorch/parallel/tensor_parallel/tensor_parallel.py:22-517`` (``apply_tensor_parallel``,
``ColumnParallelLinear``, ``RowParallelLinear``, ``VocabParallelEmbedding``) and ``tp_comms.py:229-360``
(linear with async all-reduce overlapped with the weight-gradient GEMM).

Differences by design:
* each parallel linear is ONE autograd node that owns its GEMMs *and* its collective, so the
  collective can be fused with the producing/consuming GEMM (all-gather→GEMM, GEMM→reduce-scatter
  over symmetric memory, ``scaletorch_b200.parallel.fused``) and otherwise overlaps with the wgrad
  GEMM on the c10d path;
* sequence parallelism is a property of the layers (the reference leaves the TP layers unaware of
  SP and double-reduces): with ``sequence_parallel=True`` a column layer consumes a
  token shard ``[T/tp, H]`` (all-gather in, reduce-scatter of dX in backward) and a row layer produces
  one (reduce-scatter out, all-gather of dY in backward);
* q/k/v and gate/up are fused projections, so the backward issues 2 instead of 5 collectives per layer;
* activations are token-major 2-D ``[T, H]``.
"""
from __future__ import annotations

import math

import torch
import torch.distributed as dist
import torch.nn as nn

from ... import ops
from ...ops.gemm import gemm, wgrad_into
from ..process_group import process_group_manager as pgm
from . import tp_comms as C


def _tp_defaults(group, tp_size, tp_rank):
    if tp_size is None:
        if group is not None:
            tp_size, tp_rank = dist.get_world_size(group), dist.get_rank(group)
        elif pgm:
            group, tp_size, tp_rank = pgm.tp_group, pgm.tp_world_size, pgm.tp_rank
        else:
            tp_size, tp_rank = 1, 0
    return group, tp_size, tp_rank


def _fused():
    from .. import fused as F_

    return F_


class _ColumnParallelFn(torch.autograd.Function):
    """y_local = gather?(x) · W_localᵀ ;  backward: dX = dY·W (all-reduce | reduce-scatter, overlapped
    with the wgrad GEMM), dW = dYᵀ·x_full."""

    @staticmethod
    def forward(ctx, x, weight, group, sequence_parallel, fused_comm):
        ctx.group, ctx.sp, ctx.fused = group, sequence_parallel, fused_comm
        x2 = x.reshape(-1, x.shape[-1])
        if sequence_parallel:
            F_ = _fused() if fused_comm else None
            if F_ is not None and F_.ag_gemm_available(x2, group):
                y, x_full = F_.all_gather_gemm(x2, weight, group)
            else:
                x_full = C._AllGatherDim0.forward(ctx, x2, group)
                y = gemm(x_full, weight)
            ctx.save_for_backward(x2, weight)  # keep only the shard; re-gathered in backward
            ctx.out_lead = (x_full.shape[0],)
        else:
            y = gemm(x2, weight)
            ctx.save_for_backward(x2, weight)
            ctx.out_lead = x.shape[:-1]
        ctx.x_shape = x.shape
        return y.view(*ctx.out_lead, weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        group = ctx.group
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        ws = C._ws(group)
        if ctx.sp:
            F_ = _fused() if ctx.fused else None
            # re-gather the activation shard for the wgrad GEMM (async, overlaps with dgrad)
            x_full = torch.empty((ws * x2.shape[0], x2.shape[1]), dtype=x2.dtype, device=x2.device)
            hx = dist.all_gather_into_tensor(x_full, x2.contiguous(), group=group, async_op=True)
            if F_ is not None and F_.gemm_rs_available(dy2, group):
                dx = F_.gemm_reduce_scatter(dy2, weight, group, b_mn=True)   # dgrad GEMM -> reduce-scatter, one kernel
                hr = None
            else:
                dx_full = gemm(dy2, weight, b_mn=True)
                dx = torch.empty((dx_full.shape[0] // ws, dx_full.shape[1]), dtype=dx_full.dtype, device=dx_full.device)
                if dist.get_backend(group) == "gloo":
                    dist.all_reduce(dx_full, group=group)
                    r = dist.get_rank(group)
                    dx.copy_(dx_full[r * dx.shape[0]:(r + 1) * dx.shape[0]])
                    hr = None
                else:
                    hr = dist.reduce_scatter_tensor(dx, dx_full, group=group, async_op=True)
            hx.wait()
            dw = wgrad_into(weight, dy2, x_full) if ctx.needs_input_grad[1] else None
            if hr is not None:
                hr.wait()
            return dx.view(ctx.x_shape), dw, None, None, None
        F_ = _fused() if (ctx.fused and ws > 1) else None
        if F_ is not None and F_.gemm_ar_available(dy2, weight.shape[1], group):
            dx = F_.gemm_all_reduce(dy2, weight, group, b_mn=True)
            dw = wgrad_into(weight, dy2, x2) if ctx.needs_input_grad[1] else None
            return dx.view(ctx.x_shape), dw, None, None, None
        dx = gemm(dy2, weight, b_mn=True)
        h = dist.all_reduce(dx, group=group, async_op=True) if ws > 1 else None
        dw = wgrad_into(weight, dy2, x2) if ctx.needs_input_grad[1] else None
        if h is not None:
            h.wait()
        return dx.view(ctx.x_shape), dw, None, None, None


class _RowParallelFn(torch.autograd.Function):
    """y = reduce(x_local · W_localᵀ) with reduce = all-reduce | reduce-scatter(tokens)."""

    @staticmethod
    def forward(ctx, x, weight, group, sequence_parallel, fused_comm):
        ctx.group, ctx.sp, ctx.fused = group, sequence_parallel, fused_comm
        x2 = x.reshape(-1, x.shape[-1])
        ctx.save_for_backward(x2, weight)
        ctx.x_shape = x.shape
        ws = C._ws(group)
        if sequence_parallel:
            F_ = _fused() if fused_comm else None
            if F_ is not None and F_.gemm_rs_available(x2, group):
                return F_.gemm_reduce_scatter(x2, weight, group)
            y = gemm(x2, weight)
            return C._reduce_scatter_dim0(y, group) if ws > 1 else y
        F_ = _fused() if (fused_comm and ws > 1) else None
        if F_ is not None and F_.gemm_ar_available(x2, weight.shape[0], group):
            y = F_.gemm_all_reduce(x2, weight, group)   # GEMM into the window + one-shot peer reduction
        else:
            y = gemm(x2, weight)
            if ws > 1:
                dist.all_reduce(y, group=group)
        return y.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        group = ctx.group
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        ws = C._ws(group)
        if ctx.sp and ws > 1:
            F_ = _fused() if ctx.fused else None
            if F_ is not None and F_.ag_gemm_available(dy2, group):
                dx, dy_full = F_.all_gather_gemm(dy2, weight, group, b_mn=True)
            else:
                dy_full = torch.empty((ws * dy2.shape[0], dy2.shape[1]), dtype=dy2.dtype, device=dy2.device)
                dist.all_gather_into_tensor(dy_full, dy2, group=group)
                dx = gemm(dy_full, weight, b_mn=True)
        else:
            dy_full = dy2
            dx = gemm(dy2, weight, b_mn=True)
        dw = wgrad_into(weight, dy_full, x2) if ctx.needs_input_grad[1] else None
        return dx.view(ctx.x_shape), dw, None, None, None


def _init_uniform(w: torch.Tensor, fan_in: int) -> None:
    bound = 1.0 / math.sqrt(fan_in)
    nn.init.uniform_(w, -bound, bound)


class ColumnParallelLinear(nn.Module):
    """Linear with the output dimension sharded across the TP group (weight ``[out/tp, in]``)."""

    def __init__(self, in_features: int, out_features: int, bias: bool = False, gather_output: bool = False,
                 sequence_parallel: bool = False, group=None, tp_size: int | None = None, tp_rank: int | None = None,
                 fused_comm: bool = True, init_std: float | None = None):
        super().__init__()
        self.group, self.tp_size, self.tp_rank = _tp_defaults(group, tp_size, tp_rank)
        if out_features % self.tp_size != 0:
            raise ValueError(f"out_features ({out_features}) must be divisible by tp size ({self.tp_size})")
        self.in_features, self.out_features = in_features, out_features
        self.out_features_per_partition = out_features // self.tp_size
        self.gather_output = gather_output
        self.sequence_parallel = sequence_parallel and self.tp_size > 1
        self.fused_comm = fused_comm
        self.init_std = init_std
        self.weight = nn.Parameter(torch.empty(self.out_features_per_partition, in_features))
        self.bias = nn.Parameter(torch.empty(self.out_features_per_partition)) if bias else None
        self.reset_parameters()

    def reset_parameters(self) -> None:
        if self.weight.device.type == "meta":
            return
        if self.init_std is not None:
            nn.init.normal_(self.weight, std=self.init_std)
        else:
            _init_uniform(self.weight, self.in_features)
        if self.bias is not None:
            nn.init.zeros_(self.bias)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.tp_size == 1:
            y = ops.linear(x, self.weight, self.bias)
            return y
        y = _ColumnParallelFn.apply(x, self.weight, self.group, self.sequence_parallel, self.fused_comm)
        if self.bias is not None:
            y = y + self.bias
        if self.gather_output:
            y = C.gather_from_tensor_parallel_region(y, self.group)
        return y


class RowParallelLinear(nn.Module):
    """Linear with the input dimension sharded across the TP group (weight ``[out, in/tp]``)."""

    def __init__(self, in_features: int, out_features: int, bias: bool = False, sequence_parallel: bool = False,
                 group=None, tp_size: int | None = None, tp_rank: int | None = None, fused_comm: bool = True,
                 init_std: float | None = None):
        super().__init__()
        self.group, self.tp_size, self.tp_rank = _tp_defaults(group, tp_size, tp_rank)
        if in_features % self.tp_size != 0:
            raise ValueError(f"in_features ({in_features}) must be divisible by tp size ({self.tp_size})")
        self.in_features, self.out_features = in_features, out_features
        self.in_features_per_partition = in_features // self.tp_size
        self.sequence_parallel = sequence_parallel and self.tp_size > 1
        self.fused_comm = fused_comm
        self.init_std = init_std
        self.weight = nn.Parameter(torch.empty(out_features, self.in_features_per_partition))
        self.bias = nn.Parameter(torch.empty(out_features)) if bias else None
        self.reset_parameters()

    def reset_parameters(self) -> None:
        if self.weight.device.type == "meta":
            return
        if self.init_std is not None:
            nn.init.normal_(self.weight, std=self.init_std)
        else:
            _init_uniform(self.weight, self.in_features)
        if self.bias is not None:
            nn.init.zeros_(self.bias)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if self.tp_size == 1:
            return ops.linear(x, self.weight, self.bias)
        y = _RowParallelFn.apply(x, self.weight, self.group, self.sequence_parallel, self.fused_comm)
        return y if self.bias is None else y + self.bias


class _VocabParallelEmbeddingFn(torch.autograd.Function):
    """Masked local lookup + sum over TP ranks (all-reduce, or reduce-scatter into the SP layout)."""

    @staticmethod
    def forward(ctx, ids, weight, group, vocab_start, sequence_parallel):
        ctx.group, ctx.sp, ctx.vocab_start = group, sequence_parallel, vocab_start
        flat = ids.reshape(-1)
        out = ops.embedding(flat, weight.detach(), vocab_start)
        ctx.save_for_backward(flat, weight)
        if C._ws(group) > 1:
            if sequence_parallel:
                out = C._reduce_scatter_dim0(out.contiguous(), group)
            else:
                dist.all_reduce(out, group=group)
        ctx.ids_shape = ids.shape
        return out if sequence_parallel else out.view(*ids.shape, weight.shape[1])

    @staticmethod
    def backward(ctx, dout):
        flat, weight = ctx.saved_tensors
        d2 = dout.reshape(-1, weight.shape[1]).contiguous()
        ws = C._ws(ctx.group)
        if ctx.sp and ws > 1:
            full = torch.empty((ws * d2.shape[0], d2.shape[1]), dtype=d2.dtype, device=d2.device)
            dist.all_gather_into_tensor(full, d2, group=ctx.group)
            d2 = full
        from ...ops.embedding import embedding_backward

        return None, embedding_backward(flat, d2, weight, ctx.vocab_start), None, None, None


class VocabParallelEmbedding(nn.Module):
    """Embedding with the vocabulary sharded across the TP group (reference ``tensor_parallel.py:366-507``)."""

    def __init__(self, num_embeddings: int, embedding_dim: int, padding_idx: int | None = None,
                 sequence_parallel: bool = False, group=None, tp_size: int | None = None, tp_rank: int | None = None):
        super().__init__()
        self.group, self.tp_size, self.tp_rank = _tp_defaults(group, tp_size, tp_rank)
        if num_embeddings % self.tp_size != 0:
            raise ValueError(f"num_embeddings ({num_embeddings}) must be divisible by tp size ({self.tp_size})")
        self.num_embeddings, self.embedding_dim, self.padding_idx = num_embeddings, embedding_dim, padding_idx
        self.num_embeddings_per_partition = num_embeddings // self.tp_size
        self.vocab_start_index = self.tp_rank * self.num_embeddings_per_partition
        self.vocab_end_index = self.vocab_start_index + self.num_embeddings_per_partition
        self.sequence_parallel = sequence_parallel and self.tp_size > 1
        self.weight = nn.Parameter(torch.empty(self.num_embeddings_per_partition, embedding_dim))
        self.reset_parameters()

    def reset_parameters(self) -> None:
        if self.weight.device.type != "meta":
            nn.init.normal_(self.weight, mean=0.0, std=1.0 / math.sqrt(self.embedding_dim))

    def forward(self, ids: torch.Tensor) -> torch.Tensor:
        if self.tp_size == 1:
            return ops.embedding(ids, self.weight)
        return _VocabParallelEmbeddingFn.apply(ids, self.weight, self.group, self.vocab_start_index,
                                               self.sequence_parallel)


def apply_tensor_parallel(model: nn.Module, enable_sequence_parallel: bool = False, module_mapping=None) -> nn.Module:
    """Reference-compatible entry point (``tensor_parallel.py:22-144``).

    Models of this framework are *constructed* tensor-parallel from their ``ParallelContext`` (sharded
    fused projections, vocab-parallel embedding and LM head), so for them this only validates and flips
    the sequence-parallel switch.  A foreign ``nn.Module`` following the reference's naming
    (``attention.{q,k,v}_proj / out_proj``, ``mlp.{gate,up}_proj / down_proj``, ``embedding``, ``final_proj``)
    gets its ``nn.Linear`` / embedding modules swapped for the parallel ones in place.
    """
    tp = pgm.tp_world_size if pgm else 1
    if getattr(model, "_stb_tp_constructed", False):
        if enable_sequence_parallel and tp > 1:
            model.set_sequence_parallel(True)
        return model
    if tp == 1:
        return model
    mapping = module_mapping or {
        "attention": {"q_proj": "column", "k_proj": "column", "v_proj": "column", "out_proj": "row"},
        "mlp": {"up_proj": "column", "gate_proj": "column", "down_proj": "row"},
    }

    def _swap(parent: nn.Module, name: str, kind: str):
        old = getattr(parent, name, None)
        if old is None:
            return
        if kind == "column":
            new = ColumnParallelLinear(old.in_features, old.out_features, bias=old.bias is not None,
                                       sequence_parallel=enable_sequence_parallel)
        elif kind == "row":
            new = RowParallelLinear(old.in_features, old.out_features, bias=old.bias is not None,
                                    sequence_parallel=enable_sequence_parallel)
        else:
            raise ValueError(kind)
        setattr(parent, name, new)

    for layer in getattr(model, "decoder_layers", []):
        for sub, table in mapping.items():
            mod = getattr(layer, sub, None)
            if mod is None:
                import warnings

                warnings.warn(f"layer has no '{sub}' sub-module; it stays replicated across TP ranks")
                continue
            for name, kind in table.items():
                _swap(mod, name, kind)
    emb = getattr(model, "embedding", None)
    if emb is not None and not isinstance(emb, VocabParallelEmbedding):
        model.embedding = VocabParallelEmbedding(emb.num_embeddings, emb.embedding_dim, getattr(emb, "padding_idx", None))
    fp = getattr(model, "final_proj", None)
    if fp is not None and not isinstance(fp, ColumnParallelLinear):
        model.final_proj = ColumnParallelLinear(fp.in_features, fp.out_features, bias=getattr(fp, "bias", None) is not None,
                                                gather_output=True)
    return model


def mark_tp_partial_params(model: nn.Module, tp_group, sequence_parallel: bool) -> list[str]:
    """Flag replicated parameters whose gradient is only partial on each TP rank (see
    ``FlatParamStore.sync_tp_partial_grads``); returns their names."""
    import torch.distributed as dist

    if tp_group is None or dist.get_world_size(tp_group) <= 1:
        return []
    names = []
    for n, p in model.named_parameters():
        leaf = n.rsplit(".", 2)
        is_norm = "norm" in n.split(".")[-2] if len(leaf) >= 2 else False
        if ("q_norm" in n or "k_norm" in n) or (sequence_parallel and is_norm):
            p._stb_tp_partial = True
            names.append(n)
    return names


def sync_tp_partial_grads(model: nn.Module, tp_group) -> None:
    """``.grad``-based variant for plain torch optimisers."""
    import torch.distributed as dist

    ps = [p for p in model.parameters() if getattr(p, "_stb_tp_partial", False) and p.grad is not None]
    if tp_group is None or not ps:
        return
    flat = torch.cat([p.grad.reshape(-1).float() for p in ps])
    dist.all_reduce(flat, group=tp_group)
    off = 0
    for p in ps:
        p.grad.copy_(flat[off:off + p.numel()].view_as(p.grad))
        off += p.numel()
