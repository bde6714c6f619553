"""Data parallel wrappers (reference: ``scaletorch/parallel/data_parallel/data_parallel.py:22-388``).

``DataParallelBucket`` overlaps bucketed gradient all-reduces with the backward pass; ``BasicDataParallel``
reduces per parameter.  Both reduce over ``pgm.grad_sync_group`` (CP×DP, ×EP when EP ranks hold distinct
data) for dense parameters and ``pgm.expert_grad_sync_group`` for expert parameters.

Fixed relative to the reference: with gradient accumulation the *last* micro-batch synchronises (the
reference's ``no_sync`` around all micro-steps never syncs nor exposes gradients).
"""
from __future__ import annotations

import contextlib

import torch
import torch.distributed as dist
import torch.nn as nn
from torch.autograd import Variable

from ...runtime.flat_params import FlatParamStore
from ..process_group import process_group_manager as pgm
from .bucket import BucketManager


def _groups(group, expert_group):
    if group is None and pgm:
        group = pgm.grad_sync_group
        expert_group = pgm.expert_grad_sync_group if expert_group is None else expert_group
    return group, expert_group if expert_group is not None else group


def get_store(module: nn.Module, grad_sync_group=None, fused_comm: bool = True) -> FlatParamStore:
    """Builds (once) the flat parameter / gradient store.  With a multi-rank NCCL ``grad_sync_group`` and the
    fused collectives enabled, the gradient buffer is allocated in symmetric memory (``GradWindow``)."""
    store = getattr(module, "_stb_store", None)
    if store is None:
        window = None
        p0 = next(module.parameters())
        if grad_sync_group is not None and fused_comm and p0.is_cuda:
            from ...runtime import symm

            if symm.group_ready(grad_sync_group) and symm.has_kernel("stb_allreduce_f32"):
                from ...runtime.fused_ops import GradWindow

                import os

                from ...runtime.fused_ops import ParamWindow

                holder = {}

                def alloc(numel):
                    holder["w"] = GradWindow(grad_sync_group, numel)
                    return holder["w"].grad[:numel]

                def palloc(numel, dtype):
                    holder["p"] = ParamWindow(grad_sync_group, numel, dtype)
                    return holder["p"].param[:numel]

                # the sharded optimiser needs the parameters in symmetric memory too (bf16 models only)
                sharded = (os.environ.get("SCALETORCH_B200_DP_SHARDED", "1") == "1" and p0.dtype == torch.bfloat16
                           and symm.has_kernel("stb_adamw_shard_bcast") and not (pgm and pgm.pp_world_size > 1))
                store = FlatParamStore(module, grad_alloc=alloc, param_alloc=palloc if sharded else None)
                window = holder["w"]
                store.param_window = holder.get("p")
        if store is None:
            store = FlatParamStore(module)
        store.grad_window = window
        if not hasattr(store, "param_window"):
            store.param_window = None
        module._stb_store = store
    return store


class _DPBase(nn.Module):
    def __init__(self, module: nn.Module):
        super().__init__()
        self.module = module
        self.require_backward_grad_sync = True

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def __getattr__(self, name):
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(super().__getattr__("module"), name)

    @contextlib.contextmanager
    def no_sync(self):
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old


class BasicDataParallel(_DPBase):
    """Naive DP: blocking mean all-reduce of each parameter's gradient as soon as it is accumulated."""

    def __init__(self, module: nn.Module, group=None):
        super().__init__(module)
        self.group, _ = _groups(group, None)
        self.world_size = dist.get_world_size(self.group) if self.group is not None else 1
        for p in module.parameters():
            if p.requires_grad:
                p.register_post_accumulate_grad_hook(self._hook)

    def _hook(self, p):
        if self.require_backward_grad_sync and self.world_size > 1 and p.grad is not None:
            p.grad.div_(self.world_size)
            dist.all_reduce(p.grad, group=self.group)


class DataParallelBucket(_DPBase):
    """Bucketed data parallelism over the flat gradient store: per-parameter hooks mark gradients ready, full buckets are reduced while backward continues, ``no_sync`` defers the reduction during gradient accumulation; with bf16 parameters on a peer-memory fabric the buckets are reduce-scattered for the sharded optimiser."""
    def __init__(self, module: nn.Module, bucket_size: int = 2**24, grad_type: torch.dtype | None = None,
                 group=None, expert_group=None, wire_dtype: torch.dtype | None = None):
        super().__init__(module)
        self.store = get_store(module)
        self.group, self.expert_group = _groups(group, expert_group)
        self.grad_type = grad_type or torch.float32
        self.wire_dtype = wire_dtype or torch.float32
        import os

        bucket_size = int(os.environ.get("SCALETORCH_B200_DP_BUCKET", bucket_size))
        sm_budget = int(os.environ.get("SCALETORCH_B200_DP_SM_BUDGET", 16))  # profiles/dp_allreduce_sweep.md
        self.bucket_manager = BucketManager(self.store, self.group, self.expert_group, bucket_size, self.wire_dtype,
                                            grad_window=getattr(self.store, "grad_window", None), sm_budget=sm_budget)
        self._post_backward_queued = False
        for s in self.store.slots:
            s.param.grad_ready_hook = self._make_ready_hook(s.param)
        self._install_embedding_row_exchange(module)

    def _install_embedding_row_exchange(self, module: nn.Module) -> None:
        """Embedding gradients touch at most ``tokens`` rows per rank: exchange those rows between the gradient-sync
        ranks (``runtime.fused_ops.EmbeddingGradExchange``) instead of all-reducing the ``[V, H]`` matrix, whose
        reduction would sit at the very end of backward.  Needs the peer-memory gradient window and no pipeline split
        (tied weights across stages are handled by the PP wrapper)."""
        import os

        from ...parallel.process_group import process_group_manager as pgm
        from ...parallel.tensor_parallel.tensor_parallel import VocabParallelEmbedding

        if getattr(self.store, "grad_window", None) is None or os.environ.get("SCALETORCH_B200_EMB_ROW_EXCHANGE", "1") != "1":
            return
        if pgm and pgm.pp_world_size > 1:
            return
        from ...runtime.fused_ops import EmbeddingGradExchange

        world = dist.get_world_size(self.group)
        for m in module.modules():
            if not isinstance(m, VocabParallelEmbedding):
                continue
            w = m.weight
            bucket = self.bucket_manager.param_to_bucket.get(id(w))
            if bucket is None or bucket.grad_window is None:
                continue
            tied = getattr(w, "_stb_expected_grads", 1) > 1
            if tied:
                w._stb_expected_grads -= 1        # only the LM-head wgrad arrival marks the parameter ready
                w._stb_sparse_sync = EmbeddingGradExchange(self.group, 1.0 / world, wait_for=bucket.wait_stream)
            elif len(bucket.params) == 1:
                bucket.skip_sync = True           # nothing dense to reduce for this parameter
                w._stb_sparse_sync = EmbeddingGradExchange(self.group, 1.0 / world)

    def _make_ready_hook(self, param):
        def ready(complete: bool = False):
            # a parameter used more than once per backward (tied embedding / LM head) is complete only
            # after all of its kernel-epilogue contributions arrived
            expected = getattr(param, "_stb_expected_grads", 1)
            if not complete and expected > 1:
                param._stb_arrivals = getattr(param, "_stb_arrivals", 0) + 1
                if param._stb_arrivals < expected:
                    return
                param._stb_arrivals = 0
            if not self.require_backward_grad_sync:
                return
            if not self._post_backward_queued:
                Variable._execution_engine.queue_callback(self._post_backward)
                self._post_backward_queued = True
            self.bucket_manager.mark_param_as_ready(param, sync=True)

        return ready

    def _post_backward(self) -> None:
        """Runs once at the end of a synchronising backward: launch stragglers, wait, re-arm."""
        self.bucket_manager.sync_all()      # buckets with parameters that never produced a gradient
        self.bucket_manager.wait()
        self.bucket_manager.reset()
        self._post_backward_queued = False

    def sync_grads_manually(self) -> None:
        """Reduces every bucket now (pipeline-parallel schedules call this after the last micro-batch,
        reference ``data_parallel.py:325-350``)."""
        self.bucket_manager.reset()
        self.bucket_manager.sync_all()
        self.bucket_manager.wait()
        self.bucket_manager.reset()

    def reset(self) -> None:
        """Start of a new optimisation step: forget readiness, mark gradients unwritten."""
        self.bucket_manager.reset()
        self.store.zero_grad()
        self._post_backward_queued = False
