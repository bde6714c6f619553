"""Per-iteration performance monitor (reference: ``scaletorch/utils/monitor.py:34-292``).

Differences from the reference: iteration time is measured with CUDA events on the current stream
(the reference uses ``time.time()`` without a device sync); NVML sampling also
records SM clocks and throttle reasons so benchmark numbers can be audited.
"""
from __future__ import annotations

import json
import os
import time
from collections import deque
from typing import Any

import torch

from .device import get_device_type, max_memory_allocated, memory_allocated, memory_reserved
from .logger_utils import get_logger

logger = get_logger(__name__)

MAX_HISTORY = 10_000
TRIM_TO = 5_000

try:  # optional
    import psutil
except Exception:  # pragma: no cover
    psutil = None
try:  # optional
    import pynvml
except Exception:  # pragma: no cover
    pynvml = None


class PerformanceMonitor:
    """Per-iteration timing (CUDA events, optionally resolved one step late so that logging never synchronises), throughput, memory, NVML clocks / power / throttle reasons and CPU load; bounded history, JSON dump, summary table."""
    def __init__(self, config: Any = None, log_dir: str | None = None, device_index: int | None = None):
        self.config = config
        self.log_dir = log_dir or (getattr(config, "work_dir", None) or ".")
        self.history: list[dict[str, Any]] = []
        self.start_time: float | None = None
        self._iter_t0: float | None = None
        self._ev0 = self._ev1 = None
        self._pending: deque = deque()
        self._tokens = 0
        self.total_tokens = 0
        self.iteration = 0
        self._nvml = None
        if pynvml is not None and get_device_type() == "cuda":
            try:
                pynvml.nvmlInit()
                idx = device_index if device_index is not None else torch.cuda.current_device()
                self._nvml = pynvml.nvmlDeviceGetHandleByIndex(idx)
            except Exception:
                self._nvml = None

    # ------------------------------------------------------------------ lifecycle
    def start(self) -> None:
        self.start_time = time.time()

    def start_iteration(self, tokens_processed: int = 0) -> None:
        self._tokens = tokens_processed
        self._iter_t0 = time.time()
        if get_device_type() == "cuda":
            self._ev0 = torch.cuda.Event(enable_timing=True)
            self._ev1 = torch.cuda.Event(enable_timing=True)
            self._ev0.record()

    def end_iteration(self, tokens_processed: int | None = None, sync: bool = True) -> dict[str, Any]:
        if self._iter_t0 is None:
            raise RuntimeError("end_iteration() called before start_iteration()")
        if tokens_processed is not None:
            self._tokens = tokens_processed
        if self._ev1 is not None:
            self._ev1.record()
            if sync:
                self._ev1.synchronize()
                dt = self._ev0.elapsed_time(self._ev1) / 1e3
            else:
                dt = time.time() - self._iter_t0
        else:
            dt = time.time() - self._iter_t0
        self._iter_t0 = None
        return self._record(dt, self._tokens)

    def end_iteration_deferred(self):
        """Record the end-of-step event without waiting for it; pass the handle to :meth:`resolve_iteration` later (the
        trainer resolves step ``i`` after it has enqueued step ``i + 1``, so the device queue never drains)."""
        if self._iter_t0 is None:
            raise RuntimeError("end_iteration_deferred() called before start_iteration()")
        if self._ev1 is not None:
            self._ev1.record()
        h = (self._ev0, self._ev1, self._tokens, self._iter_t0, time.time())
        self._iter_t0 = None
        return h

    def resolve_iteration(self, handle) -> dict[str, Any]:
        ev0, ev1, tokens, t0, t1 = handle
        if ev1 is not None:
            ev1.synchronize()
            dt = ev0.elapsed_time(ev1) / 1e3
        else:
            dt = t1 - t0
        return self._record(dt, tokens)

    def _record(self, dt: float, tokens: int) -> dict[str, Any]:
        self.iteration += 1
        self.total_tokens += tokens
        stats: dict[str, Any] = {
            "iteration": self.iteration,
            "iteration_time": dt,
            "tokens": tokens,
            "tokens_per_second": (tokens / dt) if dt > 0 else 0.0,
            "memory_allocated_gb": memory_allocated() / 1e9,
            "memory_reserved_gb": memory_reserved() / 1e9,
            "max_memory_allocated_gb": max_memory_allocated() / 1e9,
        }
        if psutil is not None:
            stats["cpu_percent"] = psutil.cpu_percent(interval=None)
            stats["host_memory_percent"] = psutil.virtual_memory().percent
        if self._nvml is not None:
            try:
                stats["gpu_temperature_c"] = pynvml.nvmlDeviceGetTemperature(self._nvml, pynvml.NVML_TEMPERATURE_GPU)
                stats["gpu_power_w"] = pynvml.nvmlDeviceGetPowerUsage(self._nvml) / 1e3
                stats["sm_clock_mhz"] = pynvml.nvmlDeviceGetClockInfo(self._nvml, pynvml.NVML_CLOCK_SM)
                stats["throttle_reasons"] = int(pynvml.nvmlDeviceGetCurrentClocksEventReasons(self._nvml))
            except Exception:
                pass
        self.history.append(stats)
        if len(self.history) > MAX_HISTORY:
            self.history = self.history[-TRIM_TO:]
        return stats

    # ------------------------------------------------------------------ reporting
    def get_average_stats(self, skip: int = 0) -> dict[str, Any]:
        h = self.history[skip:]
        if not h:
            return {}
        keys = [k for k, v in h[0].items() if isinstance(v, (int, float)) and k != "iteration"]
        avg = {f"avg_{k}": sum(x.get(k, 0.0) for x in h) / len(h) for k in keys}
        avg["num_iterations"] = len(h)
        avg["total_tokens"] = self.total_tokens
        if self.start_time is not None:
            avg["total_time"] = time.time() - self.start_time
        return avg

    def log_iteration_stats(self, stats: dict[str, Any]) -> None:
        logger.info("iter %d: %.1f ms, %.0f tok/s, mem %.2f GB", stats["iteration"], stats["iteration_time"] * 1e3,
                    stats["tokens_per_second"], stats["memory_allocated_gb"])

    def save_stats(self, filename: str | None = None) -> str | None:
        if not self.history:
            return None
        rank = int(os.environ.get("RANK", 0))
        filename = filename or f"performance_logs_{rank}_{time.strftime('%Y%m%d_%H%M%S')}.json"
        path = filename if os.path.isabs(filename) else os.path.join(self.log_dir, filename)
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        cfg = None
        if self.config is not None:
            try:
                cfg = {k: (v if isinstance(v, (int, float, str, bool, type(None), list, tuple)) else str(v))
                       for k, v in vars(self.config).items()}
            except TypeError:
                cfg = str(self.config)
        with open(path, "w") as f:
            json.dump({"config": cfg, "average": self.get_average_stats(), "iterations": self.history}, f, indent=1)
        return path

    def print_summary(self) -> None:
        avg = self.get_average_stats()
        if not avg:
            logger.info("no performance data recorded")
            return
        logger.info("==== performance summary (%d iterations) ====", avg["num_iterations"])
        for k in sorted(avg):
            v = avg[k]
            logger.info("  %-32s %s", k, f"{v:.4f}" if isinstance(v, float) else v)

    def close(self) -> None:
        if self._nvml is not None and pynvml is not None:
            try:
                pynvml.nvmlShutdown()
            except Exception:
                pass
            self._nvml = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
