"""One optimisation step without pipeline parallelism (reference: ``scaletorch/trainer/train_step.py:14-136``).

Differences: the loss is the fused (vocab-parallel) linear+cross-entropy of the model (no full logits);
with gradient accumulation only the last micro-batch synchronises gradients (the reference never does,
SURVEY §2.7 #1); the per-micro-batch loss stays on the device (one ``.item()`` per *step*, in the caller,
instead of one per micro-batch); the MoE auxiliary loss is added; clipping uses the global norm.
"""
from __future__ import annotations

import contextlib

import torch

from ..parallel.process_group import process_group_manager as pgm


def _micro_batch_loss(model, batch, device, gradient_checkpointing: bool, loss_scale: float, aux_loss: bool):
    input_ids = batch["input_ids"].to(device, non_blocking=True)
    target_ids = batch["target_ids"].to(device, non_blocking=True)
    raw = model.module if hasattr(model, "module") else model
    loss = raw.loss(input_ids, target_ids, gradient_checkpointing=gradient_checkpointing, loss_scale=loss_scale)
    if aux_loss:
        aux = raw.get_aux_loss()
        if aux is not None:
            loss = loss + aux * loss_scale
    return loss


def train_step(model, data_loader, device, dtype: torch.dtype = torch.bfloat16, gradient_accumulation_steps: int | None = None,
               gradient_checkpointing: bool = False, aux_loss: bool = True) -> torch.Tensor:
    """Runs ``GA`` micro-batches of forward+backward; returns the mean loss as a device scalar."""
    ga = gradient_accumulation_steps or getattr(data_loader, "gradient_accumulation_steps", 1)
    total = torch.zeros((), dtype=torch.float32, device=device)
    has_sync = hasattr(model, "no_sync")
    use_autocast = dtype == torch.bfloat16 and device.type == "cpu"   # CUDA path runs bf16 parameters natively
    for i in range(ga):
        batch = next(data_loader)
        last = i == ga - 1
        sync_ctx = contextlib.nullcontext() if (last or not has_sync) else model.no_sync()
        with sync_ctx:
            with torch.autocast("cpu", dtype=torch.bfloat16) if use_autocast else contextlib.nullcontext():
                loss = _micro_batch_loss(model, batch, device, gradient_checkpointing, 1.0 / ga, aux_loss)
            loss.backward()
        total += loss.detach().float()
    return total


def training_step(model, data_loader, device, dtype, config, tensor_shapes=None) -> torch.Tensor:
    """Schedule dispatch used by ``tools/train.py`` and ``bench.py``: AFAB / 1F1B when the model is pipeline-parallel
    (reference ``train.py:351-363``), the plain accumulate loop otherwise."""
    if pgm and pgm.pp_world_size > 1:
        from ..parallel.pipeline_parallel.pipeline_parallel import train_step_pipeline_1f1b, train_step_pipeline_afab
        from .model_builder import get_tensor_shapes

        if tensor_shapes is None:
            raw = model.module if hasattr(model, "module") else model
            tensor_shapes = get_tensor_shapes(config, raw.config)
        fn = train_step_pipeline_afab if getattr(config, "pipeline_parallel_engine", "1f1b") == "afab" else train_step_pipeline_1f1b
        return fn(model, data_loader, tensor_shapes, device, dtype, config.gradient_checkpointing)
    return train_step(model, data_loader, device, dtype, config.gradient_accumulation_steps, config.gradient_checkpointing,
                      getattr(config, "aux_loss", True))


def clip_gradients(model, max_grad_norm: float | None, optimizer=None) -> float | None:
    """Global-norm clipping.  With the fused optimiser clipping is folded into ``optimizer.step`` and this
    only reports the norm; for plain torch optimisers it clips ``.grad`` in place over the model-parallel
    group."""
    if max_grad_norm is None or max_grad_norm <= 0:
        return None
    from .optimizer import FusedAdamW

    if isinstance(optimizer, FusedAdamW):
        return None  # computed inside step(); read optimizer.grad_norm() after the step
    if pgm and pgm.tp_world_size > 1:
        from ..parallel.tensor_parallel.tensor_parallel import sync_tp_partial_grads

        sync_tp_partial_grads(model, pgm.tp_group)
    params = [p for p in model.parameters() if p.grad is not None]
    if not params:
        return 0.0
    ss = torch.stack([p.grad.detach().float().pow(2).sum() for p in params]).sum()
    if pgm and len(pgm.mp_group_ids) > 1:
        torch.distributed.all_reduce(ss, group=pgm.mp_group)
    norm = ss.sqrt()
    coef = torch.clamp(max_grad_norm / (norm + 1e-6), max=1.0)
    for p in params:
        p.grad.mul_(coef.to(p.grad.dtype))
    return float(norm.item())
