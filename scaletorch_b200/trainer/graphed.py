"""Whole-step CUDA-graph capture of the training step.

A Qwen3-0.6B step is ~650 kernel launches issued from ~900 Python autograd nodes: the host needs about as long to
issue them (≈26-30 ms) as the GPU needs to run them (≈31 ms), and with eight ranks sharing one host the step becomes
launch-bound.  Everything in the step is already shaped for capture: every kernel takes raw pointers and a stream and
allocates nothing, the gradient clip coefficient and the loss scale stay on the device, AdamW reads its
step-dependent scalars from device memory (``stb_adamw_dyn``), the peer-memory collectives advance their epochs on
the device (``csrc/common/symm.cuh``), and MoE routing never synchronises with the host.  So the whole
zero-grad → GA × (forward, backward, bucket all-reduce) → clip → AdamW sequence is captured once and replayed.

The reference has no counterpart (eager PyTorch; ``tools/optimize_mfu.py`` tries ``torch.compile``).

Captured and validated on hardware: single GPU, data parallelism (2 / 4 / 8 ranks, all-reduce and the sharded
reduce-scatter / AdamW / broadcast path), context parallelism, and -- since warm-up and capture share one side stream --
tensor parallelism (Qwen3-8B TP2 + SP: 131.8 ms replayed vs 133.7 ms eager) and expert-parallel MoE (Qwen3-30B-A3B
shapes, 8 layers, EP2: 56.1 vs 57.9 ms), 2 ranks each (gpurun_out/n2c_*.log -> profiles/multigpu_r2_n2.jsonl).
``SCALETORCH_B200_GRAPH_MP`` lists the model-parallel layouts that may be captured (default ``tp,ep``; ``none`` keeps
them eager).  Not captured: pipeline-parallel schedules (NCCL p2p with data-dependent ordering across micro-batches) and
the c10d ZeRO-1 fallback.  ``GraphedTrainStep.supported`` says whether a configuration qualifies.

Two things that were needed beyond "no host sync": (1) the eager warm-up steps run on the SAME side stream the capture
uses -- AccumulateGrad nodes remember the stream they were created on, and with warm-up on the default stream the
captured backward made the default stream wait on a captured event, which invalidated MoE captures at ``capture_end``
(``cudaErrorStreamCaptureInvalidated``) with no failing call before it; (2) graphs are released before the process
group is destroyed (``release_all``) -- a graph holding NCCL kernel nodes stalled the communicator teardown.
"""
from __future__ import annotations

import weakref

import torch

from ..parallel.process_group import process_group_manager as pgm
from .train_step import training_step

_live: "weakref.WeakSet[GraphedTrainStep]" = weakref.WeakSet()


def release_all() -> None:
    """Drops every captured step graph.  ``cleanup_distributed`` calls it before ``destroy_process_group``: a graph that
    holds NCCL kernel nodes keeps the communicator busy, and tearing the communicator down under it stalled the one
    TP2 run that exited with the graph alive."""
    for g in list(_live):
        g.release()


class _StaticBatches:
    """Iterator over the static device buffers the captured step reads (one per micro-batch)."""

    def __init__(self, buffers, grad_accum: int):
        self.buffers, self.gradient_accumulation_steps, self._i = buffers, grad_accum, 0

    def __next__(self):
        b = self.buffers[self._i % len(self.buffers)]
        self._i += 1
        return b


class GraphedTrainStep:
    """``step = GraphedTrainStep(model, optimizer, config, device); loss = step(data_loader)``.

    The first call runs ``warmup`` eager steps on a side stream (real optimiser steps on real batches), captures the next
    one, and from then on every call copies the batch into the static buffers, refreshes the optimiser scalars and
    replays the graph.  The returned loss is a device scalar owned by the graph (valid until the next call).
    """

    def __init__(self, model, optimizer, config, device, warmup: int = 3):
        self.model, self.optimizer, self.config, self.device = model, optimizer, config, device
        self.warmup = warmup
        self.graph: torch.cuda.CUDAGraph | None = None
        self.static_loss: torch.Tensor | None = None
        self._buffers: list[dict] | None = None
        self._steps_seen = 0
        self.failed = False
        self.ga = config.gradient_accumulation_steps
        import os

        # Warm-up and capture on ONE side stream (the documented recipe).  Autograd nodes that outlive an iteration --
        # the AccumulateGrad node of a parameter whose gradient arrives through plain autograd -- remember the stream
        # they were created on; if that is the default stream, the capture makes the default stream wait on a captured
        # event, which invalidates the capture (cudaErrorStreamCaptureInvalidated at capture_end).
        self._stream = torch.cuda.Stream() if os.environ.get("SCALETORCH_B200_GRAPH_SIDE_STREAM", "1") == "1" else None
        _live.add(self)

    def release(self) -> None:
        """Waits for the device, then frees the graph and the memory pool it owns (the step falls back to eager)."""
        if self.graph is not None:
            torch.cuda.synchronize()
            self.graph.reset()
            self.graph, self.static_loss = None, None
            self.failed = True                       # do not re-capture after a release

    @staticmethod
    def supported(config, optimizer, model=None) -> bool:
        from .optimizer import FusedAdamW

        if not torch.cuda.is_available() or not isinstance(optimizer, FusedAdamW):
            return False
        if hasattr(optimizer, "shards"):        # c10d ZeRO-1 broadcasts parameters through NCCL: not captured
            return False
        import os

        allow = os.environ.get("SCALETORCH_B200_GRAPH_MP", "tp,ep")  # model-parallel layouts that may be captured
        if pgm and ((pgm.ep_world_size > 1 and "ep" not in allow) or (pgm.tp_world_size > 1 and "tp" not in allow)):
            # SCALETORCH_B200_GRAPH_MP=none (or a list without the layout) keeps these eager
            return False
        if model is not None:
            from ..models.model_qwen3_moe import MoELayer

            if any(isinstance(m, MoELayer) for m in model.modules()) and "ep" not in allow:
                return False
        if pgm and pgm.pp_world_size > 1:
            return False
        if pgm and pgm.world_size > 1 and not getattr(config, "fused_comm", True):
            return False
        return True

    # ------------------------------------------------------------------------------------------------------
    def _eager(self, src) -> torch.Tensor:
        if hasattr(self.optimizer, "gather_params"):
            # sharded optimiser with a deferred parameter gather: inside a captured step the gather of the PREVIOUS
            # update sits at the beginning (communication stream, layer-wise waits), overlapping the forward pass
            self.optimizer.gather_params(force=torch.cuda.is_current_stream_capturing())
        self.optimizer.zero_grad()
        if hasattr(self.model, "reset"):
            self.model.reset()
        loss = training_step(self.model, src, self.device, torch.bfloat16, self.config)
        self.optimizer.step()
        return loss

    def _fetch(self, data_loader) -> list[dict]:
        return [next(data_loader) for _ in range(self.ga)]

    def _stage(self, batches: list[dict]) -> None:
        if self._buffers is None:
            self._buffers = [{k: torch.empty_like(v, device=self.device) for k, v in b.items() if torch.is_tensor(v)}
                             for b in batches]
        for buf, b in zip(self._buffers, batches):
            for k, dst in buf.items():
                dst.copy_(b[k], non_blocking=True)

    def __call__(self, data_loader) -> torch.Tensor:
        batches = self._fetch(data_loader)
        self._stage(batches)
        src = _StaticBatches(self._buffers, self.ga)
        if self.graph is None:
            if self._steps_seen < self.warmup:
                self._steps_seen += 1
                if self._stream is None:
                    return self._eager(src)
                cur = torch.cuda.current_stream()
                self._stream.wait_stream(cur)
                with torch.cuda.stream(self._stream):
                    loss = self._eager(src)
                cur.wait_stream(self._stream)
                return loss
            # capture; the captured work also *runs* once?  No: capture only records, so run the step for this batch by
            # replaying right after the capture.
            if self.failed:
                return self._eager(src)
            torch.cuda.synchronize()
            torch.cuda.empty_cache()          # the eager warm-up pool is not needed again: the graph owns its memory
            graph = torch.cuda.CUDAGraph()
            try:
                # thread_local: the loader's prefetch thread may pin memory while we capture; kernels launched into the
                # capturing stream by the autograd thread are captured in either mode
                with torch.cuda.graph(graph, stream=self._stream, capture_error_mode="thread_local"):
                    self.static_loss = self._eager(src)
                self.graph = graph
            except Exception as e:  # an op in this configuration cannot be captured: stay eager, loudly
                import traceback
                import warnings

                where = "".join(traceback.format_exc().splitlines(keepends=True)[-14:])
                warnings.warn(f"CUDA-graph capture of the training step failed ({type(e).__name__}: "
                              f"{str(e).splitlines()[0] if str(e) else e}); running eagerly.  Raised at:\n{where}")
                self.failed = True
                torch.cuda.synchronize()
                return self._eager(_StaticBatches(self._buffers, self.ga))
        self.optimizer.prepare_step()
        self.graph.replay()
        return self.static_loss
