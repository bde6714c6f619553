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
    """Sum of a small fp32 tensor over ``group`` (c10d; it captures into the step graph).  The sharded optimiser uses the
    peer-memory ``stb_allreduce_tiny_f32`` for its own partial sums (``trainer/zero.py``)."""
    dist.all_reduce(t, group=group)


class FusedAdamW(torch.optim.Optimizer):
    """AdamW over a ``FlatParamStore``: sum-of-squares kernel, exact global-norm clip coefficient computed on the device (reduced over ``norm_group``, replicated parameters counted once), one update kernel over the flat buffers; LR and bias corrections live in device memory so the step replays from a CUDA graph."""
    def __init__(self, store: FlatParamStore, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, max_grad_norm: float | None = None, norm_group=None,
                 norm_scale_tp_replicated: bool = True, _alloc_state: bool = True):
        self.store = store
        defaults = dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay)
        super().__init__(store.parameters(), defaults)
        if _alloc_state:
            self.exp_avg = torch.zeros(store.total_numel, dtype=torch.float32, device=store.device)
            self.exp_avg_sq = torch.zeros(store.total_numel, dtype=torch.float32, device=store.device)
        self.step_count = 0
        self.max_grad_norm = max_grad_norm
        self.norm_group = norm_group
        self.last_grad_norm: torch.Tensor | None = None  # device scalar (read lazily for logging)
        # Expert weights receive the token contributions of every EP rank, each differentiated from that rank's own
        # (un-averaged) loss, so their raw gradient is ep x the gradient of the mean loss; dense gradients get their
        # 1/ep from the mean all-reduce over the data replicas.  The factor is folded into the expert range's AdamW
        # launch and norm contribution instead of a separate pass over the expert gradients.
        from ..parallel.process_group import process_group_manager as _pgm

        ep = _pgm.ep_world_size if _pgm else 1
        self.expert_grad_scale = 1.0 / ep if (ep > 1 and store.expert_numel > 0 and norm_group is not None) else 1.0
        # step-dependent scalars {lr, 1-β1^t, 1-β2^t} live in device memory so the whole step can sit in a CUDA graph
        self._hyper_dev = self._hyper_host = None
        if store.device.type == "cuda":
            self._hyper_dev = torch.zeros(4, dtype=torch.float32, device=store.device)
            self._hyper_host = torch.zeros(4, dtype=torch.float32, pin_memory=True)

    def add_param_group(self, param_group) -> None:
        """``Optimizer.add_param_group`` without its ``torch._dynamo.disable`` wrapper: the wrapper's first call imports
        ``torch._dynamo`` (seconds on every rank of every launch) and nothing here is ever traced."""
        plain = getattr(torch.optim.Optimizer.add_param_group, "__wrapped__", None)
        if plain is None:
            super().add_param_group(param_group)
        else:
            plain(self, param_group)

    def _local_sumsq(self) -> torch.Tensor:
        """This rank's share of the global Σg²: every parameter counted exactly once across the model-parallel group.

        A parameter that is replicated on ``r`` ranks of the group (after gradient synchronisation its gradient is
        identical on all of them) contributes ``Σg² / r`` here: dense tensor-parallel shards are replicated over EP
        (r = ep), dense non-sharded parameters (norm weights, router) over TP × EP, experts over TP, and a tied
        embedding / LM head split across the first and last pipeline stage once more.  (Summing the plain local
        Σg² over the group, as a first version did, counted norm weights tp times; the reference clips with the
        rank-local norm, SURVEY §2.7 #13.)"""
        from ..parallel.process_group import process_group_manager as pgm

        store = self.store
        tp = pgm.tp_world_size if pgm else 1
        ep = pgm.ep_world_size if pgm else 1
        pp = pgm.pp_world_size if pgm else 1
        if self.norm_group is None or (tp == 1 and ep == 1 and pp == 1):   # no model parallelism: nothing is replicated
            return sum_of_squares(store.grad_flat[:store.total_numel])
        g = store.grad_flat
        ss_dense = sum_of_squares(g[:store.dense_numel])
        ss_exp = sum_of_squares(g[store.dense_numel:store.total_numel]) if store.expert_numel > 0 else None
        rep = getattr(self, "_norm_classes", None)
        if rep is None:
            from ..utils.misc import is_tp_sharded_name

            tp_rep = [s for s in store.slots if s.segment == "dense" and tp > 1 and not is_tp_sharded_name(s.name)]
            tied = []
            if pp > 1:
                root = getattr(store, "module", None)
                if root is not None and getattr(root, "tie_word_embeddings", False):
                    tied = [s for s in store.slots if s.name in ("embedding.weight", "final_proj.weight")]
            rep = self._norm_classes = (tp_rep, tied)
        tp_rep, tied = rep

        def ss_of(slots):
            if not slots:
                return None
            return torch.stack([s.param.main_grad.float().pow(2).sum() for s in slots]).sum().reshape(1)

        total = ss_dense / ep
        s_rep = ss_of(tp_rep)
        if s_rep is not None:
            total = total - s_rep * ((1.0 - 1.0 / tp) / ep)
        s_tied = ss_of(tied)
        if s_tied is not None:                      # the same matrix lives on the first and the last stage
            r_t = ep * (tp if tp_rep and tied[0] in tp_rep else 1)
            total = total - s_tied * (0.5 / r_t)
        if ss_exp is not None:
            total = total + ss_exp * (self.expert_grad_scale ** 2 / tp)
        return total

    @torch.no_grad()
    def compute_clip(self):
        """Device scalars ``(coef, norm)`` of the global gradient norm over the model-parallel group."""
        ss = self._local_sumsq()
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
        self._apply_update(g, coef, grad_scale)
        return loss

    def _apply_update(self, g: dict, coef, grad_scale: float) -> None:
        """AdamW over the flat buffers: one launch, or two when the expert range carries its own gradient scale."""
        st = self.store
        ranges = [(0, st.total_numel, grad_scale)]
        if self.expert_grad_scale != 1.0:
            ranges = [(0, st.dense_numel, grad_scale), (st.dense_numel, st.total_numel, grad_scale * self.expert_grad_scale)]
        for lo, hi, gs in ranges:
            if hi > lo:
                fused_adamw_step(st.param_flat[lo:hi], st.grad_flat[lo:hi], self.exp_avg[lo:hi], self.exp_avg_sq[lo:hi],
                                 lr=g["lr"], betas=g["betas"], eps=g["eps"], weight_decay=g["weight_decay"],
                                 step=max(self.step_count, 1), clip_coef=coef, grad_scale=gs, hyper=self._hyper_dev)

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
