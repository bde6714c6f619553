"""Learning-rate schedulers with linear warm-up and a plug-in registry.

Reference: ``scaletorch/trainer/lr_scheduler.py:27-211`` (factory for linear / cosine / polynomial / step /
onecycle via ``LambdaLR`` + ``register_scheduler``).  ``constant`` is added.
"""
from __future__ import annotations

import math
from typing import Callable

from torch.optim import Optimizer
from torch.optim.lr_scheduler import LambdaLR, LRScheduler, OneCycleLR

_REGISTRY: dict[str, Callable] = {}


def register_scheduler(name: str, factory: Callable | None = None):
    """Registers ``factory(optimizer, config, total_steps) -> LRScheduler`` (usable as a decorator)."""

    def _reg(fn: Callable) -> Callable:
        _REGISTRY[name] = fn
        return fn

    return _reg(factory) if factory is not None else _reg


def registered_schedulers() -> list[str]:
    """Names accepted by ``--lr_scheduler_type`` (built-ins plus ``register_scheduler`` plug-ins)."""
    return sorted(_REGISTRY)


def _with_warmup(decay: Callable[[int], float], warmup: int) -> Callable[[int], float]:
    def fn(step: int) -> float:
        if warmup > 0 and step < warmup:
            return float(step + 1) / float(warmup)
        return decay(step - warmup)

    return fn


@register_scheduler("constant")
def _constant(optimizer, config, total_steps):
    return LambdaLR(optimizer, _with_warmup(lambda s: 1.0, config.warmup_steps))


@register_scheduler("linear")
def _linear(optimizer, config, total_steps):
    span = max(1, total_steps - config.warmup_steps)
    return LambdaLR(optimizer, _with_warmup(lambda s: max(0.0, 1.0 - s / span), config.warmup_steps))


@register_scheduler("cosine")
def _cosine(optimizer, config, total_steps):
    t_max = config.T_max or max(1, total_steps - config.warmup_steps)
    base = optimizer.param_groups[0]["lr"]
    floor = (config.eta_min / base) if base > 0 else 0.0

    def decay(s: int) -> float:
        s = min(s, t_max)
        return floor + (1.0 - floor) * 0.5 * (1.0 + math.cos(math.pi * s / t_max))

    return LambdaLR(optimizer, _with_warmup(decay, config.warmup_steps))


@register_scheduler("polynomial")
def _polynomial(optimizer, config, total_steps):
    span = max(1, total_steps - config.warmup_steps)
    return LambdaLR(optimizer, _with_warmup(lambda s: max(0.0, 1.0 - min(s, span) / span) ** config.power,
                                            config.warmup_steps))


@register_scheduler("step")
def _step(optimizer, config, total_steps):
    return LambdaLR(optimizer, _with_warmup(lambda s: config.gamma ** (s // config.step_size), config.warmup_steps))


@register_scheduler("onecycle")
def _onecycle(optimizer, config, total_steps):
    max_lr = config.max_lr or optimizer.param_groups[0]["lr"]
    return OneCycleLR(optimizer, max_lr=max_lr, total_steps=max(total_steps, 2), pct_start=config.pct_start)


def create_lr_scheduler(optimizer: Optimizer, config, total_steps: int | None = None) -> LRScheduler:
    """Build the scheduler named by ``config.lr_scheduler_type`` for ``total_steps`` (default: ``config.total_train_steps``)."""
    total = total_steps or getattr(config, "total_train_steps", None) or 10_000
    name = config.lr_scheduler_type
    if name not in _REGISTRY:
        raise ValueError(f"unknown lr_scheduler_type '{name}'; registered: {registered_schedulers()}")
    return _REGISTRY[name](optimizer, config, int(total))
