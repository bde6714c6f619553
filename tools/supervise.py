"""Run a training command under supervision: if it dies, relaunch it from the newest complete checkpoint.

    python tools/supervise.py --work_dir work_dir --max_restarts 3 -- \
        python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/train.py … --work_dir work_dir --save_frequency 100

The reference's recovery story is "relaunch by hand with ``--resume_path``" (`SURVEY.md` §5.3: no heartbeat, no in-job
restart).  This is the small missing piece on top of what the trainer already has (step watchdog with stack dumps,
per-rank heartbeat files, atomic checkpoint publish, resume of weights / optimiser / scheduler / RNG / data position):

* the command is run as a child **process group**; on a non-zero exit (a crashed rank makes torchrun exit non-zero, the
  watchdog turns a hang into an exit) the whole group is killed, so no straggler keeps a GPU or a port;
* the newest checkpoint directory ``<work_dir>/<step>`` that holds every shard file it should (same number of
  ``weights_*.pth`` files as the first complete one -- a directory caught half-written by the crash is skipped) is appended
  as ``--resume_path`` (replacing an earlier one) and the command is started again, at most ``--max_restarts`` times;
* a restart that makes no progress (dies before writing a newer checkpoint, ``--max_stalls`` times in a row) stops the loop:
  the fault is deterministic and needs a human;
* one JSON line per attempt goes to ``<work_dir>/supervisor.jsonl`` (attempt, return code, seconds, resume step).

Pure host code, no GPU needed; ``tests/test_supervise_cpu.py`` kills a CPU run with the trainer's fault injector and
checks that the supervised job still finishes with the right number of steps.
"""
from __future__ import annotations

import argparse
import json
import os
import signal
import subprocess
import sys
import time


def complete_checkpoints(work_dir: str) -> list[tuple[int, str]]:
    """``[(step, path)]`` of checkpoint directories that hold a full set of shard files, oldest first.  Same rule as
    ``scaletorch_b200.utils.checkpoint.find_complete_checkpoints`` (``--resume_path latest``), restated here so that the
    supervisor process never imports torch; ``tests/test_supervise_cpu.py`` checks that the two agree."""
    found = []
    if not os.path.isdir(work_dir):
        return found
    for name in os.listdir(work_dir):
        path = os.path.join(work_dir, name)
        if name.isdigit() and os.path.isdir(path):
            shards = [f for f in os.listdir(path) if f.startswith("weights_") and f.endswith(".pth")]
            tmp = [f for f in os.listdir(path) if f.endswith(".tmp")]
            if shards and not tmp:
                found.append((int(name), path, len(shards)))
    if not found:
        return []
    found.sort()
    full = max(n for _, _, n in found)                         # every complete step has the same number of shard files
    return [(s, p) for s, p, n in found if n == full]


def with_resume(cmd: list[str], resume_path: str | None) -> list[str]:
    """``cmd`` with ``--resume_path <path>`` set (an existing one is replaced)."""
    out, skip = [], False
    for a in cmd:
        if skip:
            skip = False
            continue
        if a == "--resume_path":
            skip = True
            continue
        if a.startswith("--resume_path="):
            continue
        out.append(a)
    return out + (["--resume_path", resume_path] if resume_path else [])


def run_once(cmd: list[str], env: dict, grace_s: float) -> int:
    proc = subprocess.Popen(cmd, env=env, start_new_session=True)          # its own process group
    try:
        rc = proc.wait()
    except KeyboardInterrupt:
        rc = 130
    # whatever the outcome, nobody of the group survives into the next attempt
    try:
        os.killpg(proc.pid, signal.SIGTERM)
        t0 = time.time()
        while time.time() - t0 < grace_s:
            try:
                os.killpg(proc.pid, 0)
            except ProcessLookupError:
                break
            time.sleep(0.1)
        else:
            os.killpg(proc.pid, signal.SIGKILL)
    except ProcessLookupError:
        pass
    return rc


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    if "--" not in argv:
        print(__doc__)
        return 2
    cut = argv.index("--")
    ap = argparse.ArgumentParser()
    ap.add_argument("--work_dir", required=True, help="the trainer's --work_dir (where <step>/ checkpoint directories appear)")
    ap.add_argument("--max_restarts", type=int, default=3)
    ap.add_argument("--max_stalls", type=int, default=2, help="consecutive restarts without a newer checkpoint before giving up")
    ap.add_argument("--grace_s", type=float, default=10.0, help="seconds between SIGTERM and SIGKILL when cleaning up a failed attempt")
    ap.add_argument("--clear_env_on_restart", default="SCALETORCH_B200_FAULT_INJECT",
                    help="comma-separated environment variables dropped after the first attempt (one-shot fault injection)")
    a = ap.parse_args(argv[:cut])
    cmd = argv[cut + 1:]
    if not cmd:
        print("supervise.py: no command after --")
        return 2
    os.makedirs(a.work_dir, exist_ok=True)
    log = os.path.join(a.work_dir, "supervisor.jsonl")
    env = dict(os.environ)
    attempt, stalls, last_step = 0, 0, -1
    while True:
        ckpts = complete_checkpoints(a.work_dir)
        step, path = ckpts[-1] if ckpts else (None, None)
        if attempt > 0:
            stalls = stalls + 1 if (step or -1) <= last_step else 0
            if stalls >= a.max_stalls:
                print(f"[supervise] {stalls} restarts without a newer checkpoint (still at step {step}): giving up", flush=True)
                return 3
        last_step = step if step is not None else -1
        run = with_resume(cmd, path) if attempt > 0 else cmd
        if attempt > 0:
            for k in filter(None, a.clear_env_on_restart.split(",")):
                env.pop(k, None)
            print(f"[supervise] restart {attempt}/{a.max_restarts} from {'step ' + str(step) if path else 'scratch'}", flush=True)
        t0 = time.time()
        rc = run_once(run, env, a.grace_s)
        with open(log, "a") as f:
            f.write(json.dumps({"attempt": attempt, "rc": rc, "seconds": round(time.time() - t0, 1), "resumed_from": step if attempt else None}) + "\n")
        if rc == 0:
            return 0
        if rc == 130 or attempt >= a.max_restarts:
            print(f"[supervise] attempt {attempt} exited with {rc}; not restarting", flush=True)
            return rc or 1
        attempt += 1


if __name__ == "__main__":
    sys.exit(main())
