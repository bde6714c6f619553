"""Fused AdamW over a :class:`FlatParamStore` with global-norm clipping folded into the update.

Reference: ``torch.optim.AdamW(fused=True)`` (``scaletorch/trainer/model_builder.py:119-134``) followed by a
separate rank-local ``clip_grad_norm_`` with a ``.item()`` sync (``trainer/train_step.py:122-136``).  Here the
step is: sum-of-squares kernel → (all-reduce of the scalar over model-parallel ranks) → clip coefficient
on the device → one AdamW kernel over the flat buffers that reads the coefficient.  No host sync.

It subclasses ``torch.optim.Optimizer`` so LR schedulers, ``param_groups`` and ``state_dict`` behave as
users expect.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops.optim import clip_coef_from_sumsq, fused_adamw_step, sum_of_squares
from ..runtime.flat_params import FlatParamStore


def _small_all_reduce(t: torch.Tensor, group) -> None:
    """Sum of a small fp32 tensor over ``group``: c10d by default, the peer-memory kernel when opted in
    (``runtime.fused_ops.SmallAllReduce``, experimental)."""
    if t.is_cuda and t.dtype == torch.float32:
        from ..runtime.fused_ops import SmallAllReduce
        from ..runtime import symm

        if SmallAllReduce.enabled() and symm.group_ready(group):
            SmallAllReduce.for_group(group).all_reduce_(t)
            return
    dist.all_reduce(t, group=group)


class FusedAdamW(torch.optim.Optimizer):
    def __init__(self, store: FlatParamStore, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, max_grad_norm: float | None = None, norm_group=None,
                 norm_scale_tp_replicated: bool = True):
        self.store = store
        defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay)
        super().__init__(store.parameters(), defaults)
        self.exp_avg = torch.zeros(store.total_numel, dtype=torch.float32, device=store.device)
        self.exp_avg_sq = torch.zeros(store.total_numel, dtype=torch.float32, device=store.device)
        self.step_count = 0
        self.max_grad_norm = max_grad_norm
        self.norm_group = norm_group
        self.last_grad_norm: torch.Tensor | None = None  # device scalar (read lazily for logging)
        # step-dependent scalars {lr, 1-β1^t, 1-β2^t} live in device memory so the whole step can sit in a CUDA graph
        self._hyper_dev = self._hyper_host = None
        if store.device.type == "cuda":
            self._hyper_dev = torch.zeros(4, dtype=torch.float32, device=store.device)
            self._hyper_host = torch.zeros(4, dtype=torch.float32, pin_memory=True)

    @torch.no_grad()
    def compute_clip(self):
        """Device scalars ``(coef, norm)`` of the global gradient norm over the model-parallel group."""
        ss = sum_of_squares(self.store.grad_flat)
        if self.norm_group is not None and dist.is_initialized() and dist.get_world_size(self.norm_group) > 1:
            _small_all_reduce(ss, self.norm_group)
        coef, norm = clip_coef_from_sumsq(ss, self.max_grad_norm or 0.0)
        self.last_grad_norm = norm
        return coef, norm

    def prepare_step(self) -> None:
        """Advance the step counter and refresh the device-side scalars.  Called by :meth:`step` in eager mode; under
        CUDA-graph replay the caller invokes it before every ``graph.replay()`` (``trainer/graphed.py``)."""
        self.step_count += 1
        if self._hyper_dev is not None:
            g = self.param_groups[0]
            b1, b2 = g["betas"]
            self._hyper_host[0] = float(g["lr"])
            self._hyper_host[1] = 1.0 - b1 ** self.step_count
            self._hyper_host[2] = 1.0 - b2 ** self.step_count
            self._hyper_dev.copy_(self._hyper_host, non_blocking=True)

    @torch.no_grad()
    def step(self, closure=None, grad_scale: float = 1.0):
        loss = closure() if closure is not None else None
        self.store.finalize_grads()
        self.store.sync_tp_partial_grads()
        capturing = self._hyper_dev is not None and torch.cuda.is_current_stream_capturing()
        if not capturing:
            self.prepare_step()
        g = self.param_groups[0]
        coef = None
        if self.max_grad_norm is not None and self.max_grad_norm > 0:
            coef, _ = self.compute_clip()
        fused_adamw_step(self.store.param_flat, self.store.grad_flat, self.exp_avg, self.exp_avg_sq,
                         lr=g["lr"], betas=g["betas"], eps=g["eps"], weight_decay=g["weight_decay"],
                         step=max(self.step_count, 1), clip_coef=coef, grad_scale=grad_scale, hyper=self._hyper_dev)
        return loss

    def zero_grad(self, set_to_none: bool = True) -> None:
        self.store.zero_grad(set_to_zero=False)

    def grad_norm(self) -> float:
        """Host value of the last global gradient norm (synchronises)."""
        return float(self.last_grad_norm.item()) if self.last_grad_norm is not None else 0.0

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq,
                "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups]}

    def load_state_dict(self, sd) -> None:
        self.step_count = int(sd["step"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        for g, s in zip(self.param_groups, sd.get("param_groups", [])):
            g.update(s)
