"""Pipeline parallelism: stage partitioning + AFAB and 1F1B schedules.

Reference: ``scaletorch/parallel/pipeline_parallel/pipeline_parallel.py:47-671``.  Same contiguous layer split
(first ``L % pp`` stages get one extra layer), first stage owns the embedding, last stage owns
``final_norm`` + ``final_proj``; micro-batches = gradient-accumulation steps; DP gradient sync deferred to after
the last micro-batch.

Fixed / changed relative to the reference: stages carry the RoPE tables (the reference's stage forward
drops ``cos/sin`` and cannot run); stage activations run in the training dtype with the
model's fused kernels and optional recomputation; the last stage uses the fused vocab-parallel
linear+cross-entropy; receives are pre-posted so p2p overlaps compute; tied embeddings across the first
and last stage are kept in sync by all-reducing their gradients.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn

from ..process_group import get_process_group_manager
from ..process_group import process_group_manager as pgm
from .pp_comms import bidirectional_pipeline_communicate, pipeline_communicate, post_recv


def distribute_layers(num_layers: int, pp_size: int, pp_rank: int) -> list[int]:
    """Contiguous split; the first ``num_layers % pp_size`` stages get one extra layer."""
    per = [num_layers // pp_size + (1 if i < num_layers % pp_size else 0) for i in range(pp_size)]
    start = sum(per[:pp_rank])
    return list(range(start, start + per[pp_rank]))


class PipelineParallel(nn.Module):
    """Holds this stage's slice of a :class:`DecoderLM` (same attribute names as the full model so the
    checkpoint mapping and ``reset_parameters`` keep working)."""

    def __init__(self, model: nn.Module, config, pp_rank: int | None = None, pp_size: int | None = None):
        super().__init__()
        self.pp_rank = pgm.pp_rank if pp_rank is None else pp_rank
        self.pp_size = pgm.pp_world_size if pp_size is None else pp_size
        self.is_first = self.pp_rank == 0
        self.is_last = self.pp_rank == self.pp_size - 1
        self.config = self.model_config = config
        self.ctx = model.ctx
        self.layer_ids = self.layer_distribution = distribute_layers(config.num_hidden_layers, self.pp_size, self.pp_rank)
        self.hidden_size = model.hidden_size
        self.vocab_size = model.vocab_size
        self.head_dim = model.head_dim
        self.default_rope_theta = model.default_rope_theta
        self.tie_word_embeddings = model.tie_word_embeddings
        self._stb_tp_constructed = True
        self.embedding = model.embedding if self.is_first else nn.Identity()
        self.decoder_layers = nn.ModuleList([model.decoder_layers[i] for i in self.layer_ids])
        self.final_norm = model.final_norm if self.is_last else nn.Identity()
        self.final_proj = model.final_proj if self.is_last else nn.Identity()
        self.register_buffer("cos", model.cos, persistent=False)
        self.register_buffer("sin", model.sin, persistent=False)
        self._model_cls_helpers = model  # borrow run_layers / finalize / lm_loss implementations
        object.__setattr__(self, "_proto", model)
        self._pipeline_loss = None
        self.tied_across_stages = bool(self.tie_word_embeddings and self.pp_size > 1 and (self.is_first or self.is_last))
        # dist.new_group is collective over the WORLD: the (first, last) pair groups are created here, by every rank
        # (middle stages included), not lazily by the two stages that use them
        if self.tie_word_embeddings and self.pp_size > 1 and pgm and dist.is_available() and dist.is_initialized() \
                and pgm.pp_world_size == self.pp_size:
            _first_last_group()

    # nn.Module bookkeeping: do not register the prototype model as a sub-module (its unused layers would
    # otherwise be materialised and counted)
    def __getattr__(self, name):
        if name == "_model_cls_helpers":
            return object.__getattribute__(self, "_proto")
        return super().__getattr__(name)

    def __setattr__(self, name, value):
        if name == "_model_cls_helpers":
            return
        super().__setattr__(name, value)

    def reset_parameters(self) -> None:
        if self.is_first:
            self.embedding.reset_parameters()
        for layer in self.decoder_layers:
            layer.reset_parameters()
        if self.is_last:
            self.final_norm.reset_parameters()
            self.final_proj.reset_parameters()

    def _apply(self, fn, *args, **kwargs):
        cos, sin = self.cos, self.sin
        out = super()._apply(fn, *args, **kwargs)
        if cos.dtype == torch.float32 and self.cos.dtype != torch.float32:
            self.cos, self.sin = cos.to(self.cos.device), sin.to(self.sin.device)
        return out

    def get_aux_loss(self):
        total = None
        for layer in self.decoder_layers:
            moe = getattr(layer, "moe", None)
            if moe is not None and moe._aux_loss is not None:
                total = moe._aux_loss if total is None else total + moe._aux_loss
        return total

    def forward(self, input_ids=None, position_ids=None, hidden_states=None, gradient_checkpointing: bool = False):
        """First stage: ``input_ids [B,S]`` → hidden ``[B,S,H]``; other stages: hidden → hidden; the last stage
        returns the normalised hidden states (feed them to :meth:`loss_from_hidden`)."""
        proto = self._proto
        if self.is_first:
            B, S = input_ids.shape
            h = self.embedding(input_ids).reshape(-1, self.hidden_size)
        else:
            # under sequence parallelism the stage boundary carries this TP rank's token shard [B, S/tp, H]
            sp_f = self.ctx.tp_size if self.ctx.sp else 1
            B, S = hidden_states.shape[0], hidden_states.shape[1] * sp_f
            h = hidden_states.reshape(-1, self.hidden_size)
        pos_offset = self.ctx.cp_rank * S if self.ctx.cp_size > 1 else 0
        positions = None
        if self.ctx.cp_size > 1:
            from ..context_parallel import zigzag

            if zigzag.enabled():
                positions = zigzag.local_positions(S, self.ctx.cp_size, self.ctx.cp_rank, h.device).repeat(B)
        delta = None
        for layer in self.decoder_layers:
            if gradient_checkpointing and torch.is_grad_enabled():
                from torch.utils.checkpoint import checkpoint

                h, delta = checkpoint(layer, h, delta, B, S, self.cos, self.sin, pos_offset, positions, use_reentrant=False)
            else:
                h, delta = layer(h, delta, B, S, self.cos, self.sin, pos_offset, positions)
        if self.is_last:
            from ... import ops

            if delta is None:
                y = ops.rmsnorm(h, self.final_norm.weight, self.final_norm.eps)
            else:
                y = ops.add_rmsnorm(delta, h, self.final_norm.weight, self.final_norm.eps)[1]
            return y
        out = h if delta is None else h + delta
        return out.view(B, -1, self.hidden_size)

    def loss_from_hidden(self, y, target_ids, loss_scale: float = 1.0):
        proto = self._proto
        return type(proto).lm_loss(self, y, target_ids, loss_scale)

    def backward(self, input_tensor, output_tensor, output_tensor_grad):
        """Backward of one micro-batch through this stage; returns the gradient w.r.t. ``input_tensor``."""
        if input_tensor is not None:
            input_tensor.retain_grad()
        if output_tensor_grad is None:
            output_tensor_grad = torch.ones_like(output_tensor, memory_format=torch.preserve_format)
        torch.autograd.backward(output_tensor, grad_tensors=output_tensor_grad)
        return input_tensor.grad if input_tensor is not None else None


def sync_tied_embedding_grads(model: PipelineParallel) -> None:
    """With tied embeddings and pp > 1 the embedding (first stage) and LM head (last stage) are copies:
    sum their gradients so both copies take the same update (the reference zero-fills the last stage's
    head and loses the tie)."""
    raw = model.module if hasattr(model, "module") else model
    if not getattr(raw, "tied_across_stages", False):
        return
    w = raw.embedding.weight if raw.is_first else raw.final_proj.weight
    g = getattr(w, "main_grad", None)
    if g is None:
        g = w.grad
    if g is None:
        return
    group = _first_last_group()
    dist.all_reduce(g, group=group)


_FL_GROUP = None
_FL_OWNER = None


def _first_last_group():
    """Group of this rank's (first stage, last stage) pair; ``None`` on middle stages.  Collective on first use."""
    global _FL_GROUP, _FL_OWNER
    if _FL_OWNER is not get_process_group_manager():
        _FL_GROUP, _FL_OWNER = None, get_process_group_manager()
        # every rank must create every (first,last) pair group in the same order
        mine = None
        grid = pgm.grid
        dp, pp, cp, ep, tp = grid.shape
        for d in range(dp):
            for c in range(cp):
                for e in range(ep):
                    for t in range(tp):
                        ranks = [int(grid[d, 0, c, e, t]), int(grid[d, pp - 1, c, e, t])]
                        g = dist.new_group(ranks=ranks)
                        if pgm.global_rank in ranks:
                            mine = g
        _FL_GROUP = mine if mine is not None else False
    return _FL_GROUP or None


def _forward_step(model, data_loader, recv, device, dtype, ga: int, gradient_checkpointing: bool):
    """One micro-batch forward on this stage.  Returns ``(input_tensor, output_tensor, loss_or_None)``."""
    raw = model.module if hasattr(model, "module") else model
    batch = next(data_loader)
    input_tensor = recv.wait() if recv is not None else None
    out = model(input_ids=batch["input_ids"].to(device, non_blocking=True) if raw.is_first else None,
                hidden_states=input_tensor, gradient_checkpointing=gradient_checkpointing)
    loss = None
    if raw.is_last:
        loss = raw.loss_from_hidden(out, batch["target_ids"].to(device, non_blocking=True), loss_scale=1.0 / ga)
        aux = raw.get_aux_loss()
        if aux is not None:
            loss = loss + aux / ga
        out = loss
    return input_tensor, out, loss


def _finish(model, losses, device):
    raw = model.module if hasattr(model, "module") else model
    sync_tied_embedding_grads(model)
    if hasattr(model, "sync_grads_manually") and pgm.grad_sync_world_size > 1:
        model.sync_grads_manually()
    if raw.is_last and losses:
        return torch.stack([l.detach().float() for l in losses]).sum()
    return torch.zeros((), dtype=torch.float32, device=device)


def train_step_pipeline_afab(model, data_loader, tensor_shapes, device, dtype, gradient_checkpointing: bool = False):
    """All-forward-all-backward schedule (reference ``pipeline_parallel.py:338-454``)."""
    ga = data_loader.gradient_accumulation_steps
    inputs, outputs, losses = [], [], []
    sync_ctx = model.no_sync() if hasattr(model, "no_sync") else _null()
    with sync_ctx:
        for _ in range(ga):
            recv = post_recv("recv_forward", tensor_shapes, device, dtype)
            inp, out, loss = _forward_step(model, data_loader, recv, device, dtype, ga, gradient_checkpointing)
            pipeline_communicate("send_forward", tensor=out, device=device, dtype=dtype)
            inputs.append(inp)
            outputs.append(out)
            if loss is not None:
                losses.append(loss)
        for _ in range(ga):
            grad = pipeline_communicate("recv_backward", shapes=tensor_shapes, device=device, dtype=dtype)
            inp, out = inputs.pop(0), outputs.pop(0)
            in_grad = (model.module if hasattr(model, "module") else model).backward(inp, out, grad)
            pipeline_communicate("send_backward", tensor=in_grad, device=device, dtype=dtype)
    return _finish(model, losses, device)


def train_step_pipeline_1f1b(model, data_loader, tensor_shapes, device, dtype, gradient_checkpointing: bool = False):
    """One-forward-one-backward schedule (reference ``pipeline_parallel.py:457-671``): warm-up =
    ``min(pp − rank − 1, GA)`` forwards, steady 1F1B with fused send/recv, cool-down backwards."""
    ga = data_loader.gradient_accumulation_steps
    raw = model.module if hasattr(model, "module") else model
    warmup = min(pgm.pp_world_size - pgm.pp_rank - 1, ga)
    remaining = ga - warmup
    inputs, outputs, losses = [], [], []
    sync_ctx = model.no_sync() if hasattr(model, "no_sync") else _null()

    def fwd(recv):
        inp, out, loss = _forward_step(model, data_loader, recv, device, dtype, ga, gradient_checkpointing)
        if loss is not None:
            losses.append(loss)
        return inp, out

    with sync_ctx:
        for _ in range(warmup):
            inp, out = fwd(post_recv("recv_forward", tensor_shapes, device, dtype))
            pipeline_communicate("send_forward", tensor=out, device=device, dtype=dtype)
            inputs.append(inp)
            outputs.append(out)
        recv = post_recv("recv_forward", tensor_shapes, device, dtype) if remaining > 0 else None
        for i in range(remaining):
            inp, out = fwd(recv)
            grad = bidirectional_pipeline_communicate("send_fwd_recv_bwd", out, tensor_shapes, device, dtype)
            inputs.append(inp)
            outputs.append(out)
            inp0, out0 = inputs.pop(0), outputs.pop(0)
            in_grad = raw.backward(inp0, out0, grad)
            if i == remaining - 1:
                recv = None
                pipeline_communicate("send_backward", tensor=in_grad, device=device, dtype=dtype)
            else:
                got = bidirectional_pipeline_communicate("send_bwd_recv_fwd", in_grad, tensor_shapes, device, dtype)
                recv = _Ready(got) if got is not None else None
        for _ in range(warmup):
            grad = pipeline_communicate("recv_backward", shapes=tensor_shapes, device=device, dtype=dtype)
            inp0, out0 = inputs.pop(0), outputs.pop(0)
            in_grad = raw.backward(inp0, out0, grad)
            pipeline_communicate("send_backward", tensor=in_grad, device=device, dtype=dtype)
    return _finish(model, losses, device)


class _Ready:
    def __init__(self, t):
        self.t = t

    def wait(self):
        return self.t


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False
