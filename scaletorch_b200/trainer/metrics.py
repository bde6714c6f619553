"""Console / wandb metrics (reference: ``scaletorch/trainer/metrics.py:23-114``).

The console line keeps the reference's exact field layout because its benchmark scripts regex-parse it
(``scripts/benchmark_comprehensive.py:231-239``).
"""
from __future__ import annotations

from typing import Any

from ..parallel.process_group import process_group_manager as pgm
from ..utils.device import memory_reserved
from ..utils.misc import get_mfu, to_readable_format


def is_logging_rank() -> bool:
    """True on the one rank that prints the metrics line: tp = dp = cp = ep = 0 on the last pipeline stage (any rank when not distributed)."""
    if not pgm:
        return True
    return pgm.tp_rank == 0 and pgm.dp_rank == 0 and pgm.cp_rank == 0 and pgm.ep_rank == 0 and pgm.pp_is_last_stage


def format_metrics_line(step: int, loss: float, lr: float, grad_norm: float | None, global_batch_size: int,
                        tokens_per_second: float, tokens_per_second_per_gpu: float, trained_tokens: int,
                        max_tokens: int | None, mfu: float, memory_gb: float) -> str:
    """The console line ``Step | Loss | LR | GradNorm | Global batch size | Tokens/s | Tokens/s/GPU | Tokens | MFU | Memory`` in the exact format the benchmark scripts parse."""
    parts = [
        f"Step: {step:<5d}",
        f"Loss: {loss:6.4f}",
        f"LR: {lr:.2e}",
        f"GradNorm: {grad_norm:6.3f}" if grad_norm is not None else "GradNorm: n/a",
        f"Global batch size: {to_readable_format(global_batch_size):>7s}",
        f"Tokens/s: {to_readable_format(tokens_per_second):>7s}",
        f"Tokens/s/GPU: {to_readable_format(tokens_per_second_per_gpu):>7s}",
        f"Tokens: {to_readable_format(trained_tokens):>7s}" + (f"/{to_readable_format(max_tokens)}" if max_tokens else ""),
        f"MFU: {mfu:5.2f}%",
        f"Memory: {memory_gb:6.2f}GB",
    ]
    return " | ".join(parts)


def log_training_metrics(config: Any, model_config: Any, step: int, loss: float, lr: float, grad_norm: float | None,
                         step_duration: float, tokens_per_step: int, trained_tokens: int, num_params: int,
                         world_size: int, logger=None, wandb_run=None) -> dict[str, Any]:
    """Compute throughput / MFU / memory for one step, print the metrics line on the logging rank, forward the dict to wandb when given, and return it."""
    tps = tokens_per_step / step_duration if step_duration > 0 else 0.0
    tps_gpu = tps / max(world_size, 1)
    mfu = get_mfu(tps_gpu, num_params, model_config, seq_len=config.sequence_length)
    mem = memory_reserved() / 1e9
    metrics = {"step": step, "loss": loss, "lr": lr, "grad_norm": grad_norm, "tokens_per_second": tps,
               "tokens_per_second_per_gpu": tps_gpu, "trained_tokens": trained_tokens, "mfu": mfu,
               "memory_gb": mem, "step_duration": step_duration}
    if is_logging_rank():
        line = format_metrics_line(step, loss, lr, grad_norm, config.global_batch_size_token or tokens_per_step, tps,
                                   tps_gpu, trained_tokens, config.max_tokens, mfu, mem)
        # the logging rank is the last PP stage, which is not global rank 0 under PP: its logger is muted
        if logger is not None and logger.isEnabledFor(20):
            logger.info(line)
        else:
            print(line, flush=True)
        if wandb_run is not None:
            wandb_run.log(metrics, step=step)
    return metrics


class AsyncScalarReader:
    """Device→host readback of per-step scalars (loss, grad norm) without draining the GPU queue.

    ``push(t)`` enqueues a copy of the device scalar into a pinned host slot followed by an event; ``pop()`` returns the
    oldest pushed value, waiting only for *its* event.  With ``lag = 1`` the training loop reads step ``i``'s loss after
    it has already enqueued step ``i + 1``, so the host keeps running ahead of the device (the reference calls
    ``.item()`` on every micro-batch, ``trainer/train_step.py:113``).  On CPU tensors it degrades to plain reads.
    """

    def __init__(self, slots: int = 4):
        import torch

        self._torch = torch
        self._cuda = torch.cuda.is_available()
        self._host = torch.zeros(slots, dtype=torch.float32, pin_memory=self._cuda)
        self._events = [torch.cuda.Event() for _ in range(slots)] if self._cuda else [None] * slots
        self._head = self._tail = 0
        self._slots = slots

    def __len__(self) -> int:
        return self._head - self._tail

    def push(self, t) -> None:
        if len(self) == self._slots:
            raise RuntimeError("AsyncScalarReader is full: pop() before pushing more")
        i = self._head % self._slots
        if t.is_cuda:
            self._host[i:i + 1].copy_(t.detach().reshape(1).float(), non_blocking=True)
            self._events[i].record()
        else:
            self._host[i] = float(t)
        self._head += 1

    def pop(self) -> float:
        if len(self) == 0:
            raise RuntimeError("AsyncScalarReader is empty")
        i = self._tail % self._slots
        if self._events[i] is not None and self._cuda:
            self._events[i].synchronize()
        self._tail += 1
        return float(self._host[i])

    def drain(self) -> list[float]:
        return [self.pop() for _ in range(len(self))]
