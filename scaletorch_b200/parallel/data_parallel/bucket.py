"""Gradient buckets as contiguous ranges of the flat fp32 gradient buffer.

Reference: ``scaletorch/parallel/data_parallel/bucket.py:22-431`` (``Bucket``, ``BucketManager``).  Kept: the
bookkeeping API (``mark_param_as_ready``, ``sync_gradient``, ``wait``, ``reset``, double-sync / double-ready
assertions) and ``bucket_size`` counted in elements (default 2**24).  Changed: buckets do not own
memory -- they are ``[start, end)`` ranges of :class:`FlatParamStore.grad_flat`, filled in reverse
registration order (≈ backward order) so they complete progressively during backward; the reduction
divides by the group size *inside* the collective launch (pre-scale) and can put bf16 on the wire.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ... import _native as N
from ...runtime.flat_params import FlatParamStore, ParamSlot


class Bucket:
    """A contiguous range of the flat gradient buffer with the parameters that live in it: tracks which gradients are ready and launches the range's reduction (peer-memory kernel on the comm stream, or c10d) when the last one arrives."""
    def __init__(self, slots: list[ParamSlot], grad_flat: torch.Tensor, group, wire_dtype: torch.dtype = torch.float32,
                 grad_window=None, comm_stream=None, sm_budget: int = 32):
        self.slots = slots
        self.params = [s.param for s in slots]
        self.start = min(s.offset for s in slots)
        self.end = max(s.offset + s.numel for s in slots)
        self.grad_data = grad_flat[self.start:self.end]
        self.group = group
        self.world_size = dist.get_world_size(group) if (group is not None and dist.is_initialized()) else 1
        self.wire_dtype = wire_dtype
        self.params_with_grad_ready: set[int] = set()
        self.handle = None
        self.launched = False          # sync_gradient() ran in this backward (the handle may already have been waited on)
        self.skip_sync = False         # every parameter of the bucket synchronises by itself (embedding row exchange)
        self._wire: torch.Tensor | None = None
        # fused path: the gradient buffer lives in symmetric memory and the bucket is reduced in place by
        # stb_allreduce_f32 (two-shot over NVLink peer memory / NVLS, 1/world scale fused) on a side stream
        self.grad_window = grad_window
        self.comm_stream = comm_stream
        self.sm_budget = sm_budget
        # set by the sharded optimiser (trainer/zero.py::PeerShardedAdamW): this rank keeps only the reduced gradient
        # of its own shard [own_lo, own_hi) of the flat buffer -> reduce-scatter instead of all-reduce
        self.own_range: tuple[int, int] | None = None

    @property
    def numel(self) -> int:
        return self.end - self.start

    def sync_gradient(self) -> None:
        """Launches the (asynchronous) mean all-reduce of this bucket."""
        if self.handle is not None or self.launched:
            raise RuntimeError("bucket gradient sync launched twice")
        self.launched = True
        if self.world_size == 1 or self.skip_sync:
            return
        for p in self.params:            # parameters that got no gradient contribute zeros
            if not p.grad_written:
                p.main_grad.zero_()
                p.grad_written = True
        inv = 1.0 / self.world_size
        if self.grad_window is not None:
            cur = torch.cuda.current_stream()
            ready = torch.cuda.Event()
            ready.record(cur)
            self.comm_stream.wait_event(ready)
            with torch.cuda.stream(self.comm_stream):
                if self.own_range is not None:
                    self.grad_window.reduce_scatter_range_(self.start, self.numel, inv, self.own_range[0],
                                                           self.own_range[1], self.sm_budget)
                else:
                    self.grad_window.all_reduce_range_(self.start, self.numel, inv, self.sm_budget)
                done = torch.cuda.Event()
                done.record(self.comm_stream)
            self.handle = _EventHandle(done)
            return
        if self.wire_dtype == torch.bfloat16 and self.grad_data.is_cuda:
            if self._wire is None:
                self._wire = torch.empty(self.numel, dtype=torch.bfloat16, device=self.grad_data.device)
            N.call("stb_cast_scale_f32_bf16", self.grad_data.data_ptr(), self._wire.data_ptr(), self.numel, inv,
                   N.stream_ptr())
            self.handle = dist.all_reduce(self._wire, group=self.group, async_op=True)
        else:
            self.grad_data.mul_(inv)
            self.handle = dist.all_reduce(self.grad_data, group=self.group, async_op=True)

    def wait(self) -> None:
        if self.handle is None:
            return
        self.handle.wait()
        self.handle = None
        if self._wire is not None and self.wire_dtype == torch.bfloat16 and self.world_size > 1:
            self.grad_data.copy_(self._wire)

    def wait_stream(self) -> None:
        """Order the current stream after this bucket's all-reduce if one is in flight (does not consume the handle)."""
        if isinstance(self.handle, _EventHandle):
            self.handle.wait()

    def reset(self) -> None:
        self.handle = None
        self.launched = False
        self.params_with_grad_ready.clear()

    def mark_param_as_ready(self, param, sync: bool = True) -> None:
        pid = id(param)
        if pid in self.params_with_grad_ready:
            raise RuntimeError("parameter marked ready twice in one backward pass")
        self.params_with_grad_ready.add(pid)
        if sync and len(self.params_with_grad_ready) == len(self.params):
            self.sync_gradient()

    def all_ready(self) -> bool:
        return len(self.params_with_grad_ready) == len(self.params)


class _EventHandle:
    """``wait()`` makes the current stream wait for the side-stream all-reduce (no host blocking)."""

    def __init__(self, event):
        self.event = event

    def wait(self):
        torch.cuda.current_stream().wait_event(self.event)


class BucketManager:
    """Partitions the store's parameters into buckets of at most ``bucket_size`` elements per reduction
    domain (dense / expert).  Oversize parameters get a bucket of their own."""

    def __init__(self, store: FlatParamStore, dense_group, expert_group=None, bucket_size: int = 2**24,
                 wire_dtype: torch.dtype = torch.float32, grad_window=None, sm_budget: int = 32):
        self.store = store
        self.grad_window = grad_window
        self.sm_budget = sm_budget
        self.comm_stream = torch.cuda.Stream() if grad_window is not None else None
        self.bucket_size = int(bucket_size)
        self.buckets: list[Bucket] = []
        self.param_to_bucket: dict[int, Bucket] = {}
        for seg, group in (("dense", dense_group), ("expert", expert_group if expert_group is not None else dense_group)):
            slots = [s for s in store.slots if s.segment == seg]
            cur: list[ParamSlot] = []
            cur_n = 0
            for s in reversed(slots):
                if cur and cur_n + s.numel > self.bucket_size:
                    self._add(cur, group, wire_dtype, seg)
                    cur, cur_n = [], 0
                cur.append(s)
                cur_n += s.numel
            if cur:
                self._add(cur, group, wire_dtype, seg)

    def _add(self, slots, group, wire_dtype, seg="dense") -> None:
        gw = self.grad_window if seg == "dense" else None   # the window is registered with the dense sync group
        b = Bucket(slots, self.store.grad_flat, group, wire_dtype, gw, self.comm_stream, self.sm_budget)
        self.buckets.append(b)
        for s in slots:
            self.param_to_bucket[id(s.param)] = b

    def mark_param_as_ready(self, param, sync: bool = True) -> None:
        self.param_to_bucket[id(param)].mark_param_as_ready(param, sync)

    def wait(self) -> None:
        for b in self.buckets:
            b.wait()

    def reset(self) -> None:
        for b in self.buckets:
            b.reset()

    def sync_all(self) -> None:
        """Launches the reduction of every bucket that has not been launched yet."""
        for b in self.buckets:
            if not b.launched:
                b.sync_gradient()
