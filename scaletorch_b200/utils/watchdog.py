"""Failure detection for training jobs: a step watchdog, a per-rank heartbeat file and a fault-injection hook.

The reference has no failure detection beyond the 180 s ``init_process_group`` timeout: a rank stuck in
a collective hangs the job until someone notices.  Kernels in this framework additionally spin on flags written by
other GPUs, so a dead peer would leave the survivors spinning; the watchdog turns that into a prompt, diagnosable exit.

* ``StepWatchdog(timeout_s)``: a daemon thread; ``beat(step)`` after every step.  If no beat arrives within the timeout it
  dumps every thread's Python stack (``faulthandler``), writes ``<dir>/hang_rank<r>.txt`` and terminates the process with
  exit code 124, so ``torchrun --max-restarts=0`` tears the job down and the launcher reports which node failed.
* Heartbeat: ``<dir>/heartbeat_rank<r>.json`` with step, wall time and pid, rewritten atomically on every beat -- an
  external supervisor (or ``scripts/torch_dist/launch_multi_nodes.sh``) can tell a slow job from a dead one.
* ``maybe_inject_fault(rank, step)``: ``SCALETORCH_B200_FAULT_INJECT="<rank>:<step>[:hang|raise|exit]"`` makes that rank
  misbehave at that step, to exercise the watchdog / clean-up / resume path in tests.
"""
from __future__ import annotations

import faulthandler
import json
import os
import sys
import threading
import time


class StepWatchdog:
    """Background thread that expects ``beat(step)`` at least every ``timeout_s`` seconds: writes a per-rank heartbeat file on every beat and, when the beats stop, dumps all thread stacks to ``hang_rank<r>.txt`` and exits the process with code 124 (or calls ``on_timeout``)."""
    def __init__(self, timeout_s: float, out_dir: str = ".", rank: int = 0, on_timeout=None, poll_s: float | None = None):
        self.timeout_s, self.out_dir, self.rank = float(timeout_s), out_dir, rank
        self.on_timeout = on_timeout
        self._last = time.monotonic()
        self._step = -1
        self._stop = threading.Event()
        self._poll = poll_s if poll_s is not None else max(0.05, min(5.0, self.timeout_s / 4))
        self._thread: threading.Thread | None = None
        self.fired = False

    def start(self) -> "StepWatchdog":
        if self.timeout_s > 0 and self._thread is None:
            os.makedirs(self.out_dir, exist_ok=True)
            self._last = time.monotonic()
            self._thread = threading.Thread(target=self._run, name="stb-watchdog", daemon=True)
            self._thread.start()
        return self

    def beat(self, step: int) -> None:
        self._last, self._step = time.monotonic(), step
        if self.timeout_s > 0:
            tmp = os.path.join(self.out_dir, f".heartbeat_rank{self.rank}.tmp")
            with open(tmp, "w") as f:
                json.dump({"step": step, "time": time.time(), "pid": os.getpid()}, f)
            os.replace(tmp, os.path.join(self.out_dir, f"heartbeat_rank{self.rank}.json"))

    def stop(self) -> None:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2 * self._poll + 1)
            self._thread = None

    def _run(self) -> None:
        while not self._stop.wait(self._poll):
            idle = time.monotonic() - self._last
            if idle > self.timeout_s:
                self.fired = True
                path = os.path.join(self.out_dir, f"hang_rank{self.rank}.txt")
                with open(path, "w") as f:
                    f.write(f"rank {self.rank}: no step completed for {idle:.1f} s after step {self._step}\n")
                    faulthandler.dump_traceback(file=f, all_threads=True)
                sys.stderr.write(f"[watchdog] rank {self.rank} stuck after step {self._step} ({idle:.0f} s); stacks in {path}\n")
                sys.stderr.flush()
                if self.on_timeout is not None:
                    self.on_timeout(self)
                else:
                    os._exit(124)
                return


def maybe_inject_fault(rank: int, step: int) -> None:
    """Fault drill: when ``SCALETORCH_B200_FAULT_INJECT=<rank>:<step>[:raise|hang|exit]`` names this rank and step, raise, sleep forever or exit with code 17."""
    spec = os.environ.get("SCALETORCH_B200_FAULT_INJECT")
    if not spec:
        return
    parts = spec.split(":")
    if int(parts[0]) != rank or int(parts[1]) != step:
        return
    mode = parts[2] if len(parts) > 2 else "raise"
    if mode == "hang":
        while True:
            time.sleep(3600)
    if mode == "exit":
        os._exit(17)
    raise RuntimeError(f"injected fault on rank {rank} at step {step}")
