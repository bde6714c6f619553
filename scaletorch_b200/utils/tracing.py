"""Tracing / profiling helpers: NVTX ranges, per-phase CUDA-event timers, and a ``torch.profiler`` window.

The reference has none of this (no ``torch.profiler``, NVTX or ``record_function`` anywhere; only the
wall-clock ``PerformanceMonitor``).  Everything here is off by default and costs nothing when off.

* ``SCALETORCH_B200_NVTX=1`` (or ``enable_nvtx()``): ``nvtx_range("fwd")`` pushes real NVTX ranges, so Nsight Systems
  timelines are annotated with step / micro-batch / phase names.
* ``PhaseTimer``: CUDA events around named phases, resolved lazily (no synchronisation inside the step).
* ``ProfilerWindow``: runs ``torch.profiler`` for steps ``[start, end)`` and writes a Chrome trace plus a kernel table.
"""
from __future__ import annotations

import contextlib
import os
from collections import defaultdict

import torch

_nvtx = os.environ.get("SCALETORCH_B200_NVTX", "0") == "1"


def enable_nvtx(on: bool = True) -> None:
    """Turn the ``nvtx_range`` annotations on or off at run time (initial value: ``SCALETORCH_B200_NVTX``)."""
    global _nvtx
    _nvtx = on


@contextlib.contextmanager
def nvtx_range(name: str):
    """Context manager: an NVTX range around the block when annotations are enabled and CUDA is present, otherwise nothing."""
    if _nvtx and torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


class PhaseTimer:
    """``with timer.phase("bwd"): ...`` records two CUDA events (wall clock on CPU); ``summary()`` synchronises once and
    returns ``{phase: {"ms_total", "ms_avg", "count"}}``."""

    def __init__(self):
        self._cuda = torch.cuda.is_available()
        self._pending: list[tuple[str, object, object]] = []
        self._acc: dict[str, list[float]] = defaultdict(list)

    @contextlib.contextmanager
    def phase(self, name: str):
        with nvtx_range(name):
            if self._cuda:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                try:
                    yield
                finally:
                    e1.record()
                    self._pending.append((name, e0, e1))
            else:
                import time

                t0 = time.perf_counter()
                try:
                    yield
                finally:
                    self._acc[name].append((time.perf_counter() - t0) * 1e3)

    def resolve(self) -> None:
        if self._pending:
            torch.cuda.synchronize()
            for name, e0, e1 in self._pending:
                self._acc[name].append(e0.elapsed_time(e1))
            self._pending.clear()

    def summary(self, reset: bool = False) -> dict[str, dict[str, float]]:
        self.resolve()
        out = {k: {"ms_total": sum(v), "ms_avg": sum(v) / len(v), "count": len(v)} for k, v in self._acc.items() if v}
        if reset:
            self._acc.clear()
        return out


class ProfilerWindow:
    """``torch.profiler`` over optimisation steps ``[start, end)``; call ``step(i)`` once per step *before* the work of
    step ``i``.  Writes ``<dir>/trace_rank<r>.json`` (Chrome / Perfetto) and ``<dir>/kernels_rank<r>.txt``."""

    def __init__(self, start: int | None, end: int | None, out_dir: str = "traces", rank: int = 0):
        self.start, self.end, self.out_dir, self.rank = start, end, out_dir, rank
        self._prof = None

    @property
    def enabled(self) -> bool:
        return self.start is not None and self.end is not None and self.end > self.start

    def step(self, i: int) -> None:
        if not self.enabled:
            return
        if i == self.start and self._prof is None:
            from torch.profiler import ProfilerActivity, profile

            acts = [ProfilerActivity.CPU] + ([ProfilerActivity.CUDA] if torch.cuda.is_available() else [])
            self._prof = profile(activities=acts, record_shapes=False)
            self._prof.__enter__()
        elif i == self.end and self._prof is not None:
            self.close()

    def close(self) -> None:
        if self._prof is None:
            return
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        self._prof.__exit__(None, None, None)
        os.makedirs(self.out_dir, exist_ok=True)
        self._prof.export_chrome_trace(os.path.join(self.out_dir, f"trace_rank{self.rank}.json"))
        key = "cuda_time_total" if torch.cuda.is_available() else "cpu_time_total"
        with open(os.path.join(self.out_dir, f"kernels_rank{self.rank}.txt"), "w") as f:
            f.write(self._prof.key_averages().table(sort_by=key, row_limit=60, max_name_column_width=100))
        self._prof = None
