"""Auxiliary subsystems: tracing helpers, watchdog / heartbeat / fault injection, async scalar read-back."""
import json
import os
import time

import pytest
import torch

from scaletorch_b200.trainer import AsyncScalarReader
from scaletorch_b200.utils.tracing import PhaseTimer, ProfilerWindow, nvtx_range
from scaletorch_b200.utils.watchdog import StepWatchdog, maybe_inject_fault


def test_phase_timer_and_nvtx_are_noops_on_cpu():
    t = PhaseTimer()
    for _ in range(3):
        with t.phase("fwd"):
            torch.randn(64, 64) @ torch.randn(64, 64)
        with nvtx_range("x"), t.phase("bwd"):
            pass
    s = t.summary()
    assert s["fwd"]["count"] == 3 and s["bwd"]["count"] == 3 and s["fwd"]["ms_total"] >= 0


def test_profiler_window_writes_trace(tmp_path):
    w = ProfilerWindow(1, 3, str(tmp_path), rank=0)
    for i in range(5):
        w.step(i)
        torch.randn(32, 32) @ torch.randn(32, 32)
    w.close()
    assert (tmp_path / "trace_rank0.json").exists() and (tmp_path / "kernels_rank0.txt").exists()
    assert not ProfilerWindow(None, None).enabled


def test_watchdog_heartbeat_and_timeout(tmp_path):
    fired = []
    wd = StepWatchdog(0.3, str(tmp_path), rank=3, on_timeout=lambda w: fired.append(w._step), poll_s=0.05).start()
    for s in range(3):
        time.sleep(0.05)
        wd.beat(s)
    hb = json.load(open(tmp_path / "heartbeat_rank3.json"))
    assert hb["step"] == 2 and hb["pid"] == os.getpid()
    assert not wd.fired
    time.sleep(0.6)                      # no beats: the watchdog must fire and dump stacks
    assert wd.fired and fired == [2]
    assert "no step completed" in open(tmp_path / "hang_rank3.txt").read()
    wd.stop()


def test_fault_injection(monkeypatch):
    monkeypatch.setenv("SCALETORCH_B200_FAULT_INJECT", "1:5")
    maybe_inject_fault(0, 5)
    maybe_inject_fault(1, 4)
    with pytest.raises(RuntimeError, match="injected fault"):
        maybe_inject_fault(1, 5)


def test_async_scalar_reader_cpu():
    r = AsyncScalarReader(slots=3)
    r.push(torch.tensor(1.5))
    r.push(torch.tensor([2.5]))
    assert len(r) == 2 and r.pop() == 1.5
    r.push(torch.tensor(3.0))
    r.push(torch.tensor(4.0))
    with pytest.raises(RuntimeError):
        r.push(torch.tensor(5.0))
    assert r.drain() == [2.5, 3.0, 4.0]
