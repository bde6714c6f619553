"""Held-out loss: the pass that ``tools/evaluate.py`` runs once and ``tools/train.py --eval_interval N`` runs every N steps.

The reference has no evaluation path at all (its trainer only logs the training loss, ``tools/train.py:371-392``).  What a
user switching over still needs is small: the same ``model.loss`` the training step calls, under ``no_grad`` and in
``eval()`` mode, over a FIXED set of held-out windows, with token counts and loss sums added up over the ranks that hold
different tokens.  Three rules keep the number meaningful:

* every evaluation reads the same batches (the loader is rewound first), so the curve is comparable step to step;
* a token is counted once: data-parallel ranks hold different windows, context-parallel ranks different chunks of the
  same windows, tensor-parallel ranks identical values -- the sum runs over the gradient-synchronisation group
  (CP x DP x EP), exactly the ranks whose training losses are averaged;
* evaluation does not perturb training: it has its own loader (the training stream's position is untouched), draws no
  random numbers, and leaves the model in ``train()`` mode.
"""
from __future__ import annotations

import math
import time
from typing import Any

import torch

from .. import dist as D
from ..parallel.process_group import process_group_manager as pgm


@torch.no_grad()
def evaluate_loss(model, loader, device, max_batches: int | None = None, tensor_shapes=None, dtype=None) -> dict:
    """Sum of token losses and token count over this rank's next ``max_batches`` batches (default: one epoch).

    With pipeline stages (``tensor_shapes`` = the activation shape the stages exchange, as in training) the pass is
    forward-only through the pipeline: every stage receives, runs its layers and sends on; the last stage holds the loss
    and the token count, the other stages report zeros."""
    raw = model.module if hasattr(model, "module") else model
    while hasattr(raw, "module") and isinstance(raw.module, torch.nn.Module):
        raw = raw.module
    was_training = raw.training
    raw.eval()
    pipelined = bool(pgm) and pgm.pp_world_size > 1
    if pipelined:
        from ..parallel.pipeline_parallel.pp_comms import pipeline_communicate, post_recv

        if tensor_shapes is None or dtype is None:
            raise ValueError("evaluation under pipeline parallelism needs tensor_shapes and dtype")
    try:
        n_batches = len(loader) if max_batches is None else min(len(loader), max_batches)
        loss_sum = torch.zeros((), dtype=torch.float64, device=device)
        tokens = 0
        t0 = time.time()
        for _ in range(n_batches):
            b = next(loader)
            x, y = b["input_ids"].to(device, non_blocking=True), b["target_ids"].to(device, non_blocking=True)
            if pipelined:
                recv = post_recv("recv_forward", tensor_shapes, device, dtype)
                hidden = recv.wait() if recv is not None else None
                out = raw(input_ids=x if raw.is_first else None, hidden_states=hidden)
                if raw.is_last:
                    n = int((y != -100).sum())
                    loss_sum += raw.loss_from_hidden(out, y, loss_scale=1.0).double() * n
                    tokens += n
                else:
                    pipeline_communicate("send_forward", tensor=out, device=device, dtype=dtype)
                continue
            n = int((y != -100).sum())
            loss_sum += raw.loss(x, y).double() * n          # mean over the batch's tokens -> sum
            tokens += n
        if device.type == "cuda":
            torch.cuda.synchronize()
        return {"loss_sum": loss_sum, "tokens": tokens, "batches": n_batches, "seconds": time.time() - t0}
    finally:
        raw.train(was_training)


def reduce_eval_stats(part: dict, device) -> dict:
    """Combine the per-rank sums of :func:`evaluate_loss` into ``{"loss", "perplexity", "tokens", "batches", "seconds"}``."""
    stats = {"loss_sum": part["loss_sum"].reshape(1),
             "tokens": torch.tensor([float(part["tokens"])], dtype=torch.float64, device=device)}
    if pgm and pgm.grad_sync_world_size > 1:
        D.all_reduce_dict(stats, op="sum", group=pgm.grad_sync_group)
    tokens = float(stats["tokens"].item())
    loss = float(stats["loss_sum"].item()) / max(tokens, 1.0)
    return {"loss": loss, "perplexity": math.exp(min(loss, 50.0)), "tokens": int(tokens), "batches": int(part["batches"]),
            "seconds": float(part["seconds"])}


class PeriodicEvaluator:
    """Every ``interval`` optimiser steps: rewind the held-out loader, run :func:`evaluate_loss` over ``batches`` batches per
    rank, reduce, and hand the result to ``report(step, result)``.  ``interval <= 0`` disables it."""

    def __init__(self, model, loader, device, interval: int, batches: int, report=None, tensor_shapes=None, dtype=None):
        self.model, self.loader, self.device = model, loader, device
        self.tensor_shapes, self.dtype = tensor_shapes, dtype
        self.interval, self.batches, self.report = int(interval), int(batches), report
        self.history: list[tuple[int, dict[str, Any]]] = []

    def due(self, step: int) -> bool:
        return self.interval > 0 and self.loader is not None and step > 0 and step % self.interval == 0

    def run(self, step: int) -> dict:
        if self.device.type == "cuda":
            torch.cuda.synchronize()                        # the step (possibly a graph replay on a side stream) is complete
        self.loader.load_state_dict({"batches_served": 0})   # same held-out batches every time
        result = reduce_eval_stats(evaluate_loss(self.model, self.loader, self.device, self.batches, self.tensor_shapes,
                                                 self.dtype), self.device)
        self.history.append((step, result))
        if self.report is not None:
            self.report(step, result)
        return result

    def maybe_run(self, step: int) -> dict | None:
        return self.run(step) if self.due(step) else None


def format_eval_line(step: int, result: dict) -> str:
    """Console line of one evaluation, in the style of the training metrics line."""
    return (f"Eval | Step: {step:<5d} | Loss: {result['loss']:6.4f} | Perplexity: {result['perplexity']:10.3f} | "
            f"Tokens: {result['tokens']} | Batches/rank: {result['batches']} | Time: {result['seconds']:.2f}s")
