"""``tools/train.py --save_on_signal true``: SIGTERM mid-run makes every rank stop after the same step, write a
checkpoint and leave cleanly; the job can be resumed from it."""
import os
import re
import signal
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _start(args, nproc=1):
    base = [sys.executable, "-u"]
    if nproc > 1:
        base = [sys.executable, "-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1", "--nproc-per-node", str(nproc)]
    cmd = base + [os.path.join(ROOT, "tools", "train.py"), "--synthetic_data", "true", "--use_cpu", "true", "--model_name_or_path", "tiny-llama",
                  "--sequence_length", "64", "--micro_batch_size", "2", "--learning_rate", "1e-3", "--log_lag", "0",
                  "--lr_scheduler_type", "step", "--step_size", "100000", "--save_on_signal", "true", "--save_frequency", "0"] + args
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2", PYTHONUNBUFFERED="1")
    return subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, cwd=ROOT, start_new_session=True)


def _wait_for_step(proc, n, timeout=300):
    """Reads the child's output until ``Step: n`` has been logged; returns what was read."""
    import select

    buf, t0 = "", time.time()
    while time.time() - t0 < timeout:
        ready, _, _ = select.select([proc.stdout], [], [], 1.0)   # never block for good on a silent child
        if not ready:
            if proc.poll() is not None:
                break
            continue
        line = proc.stdout.readline()
        if not line:
            if proc.poll() is not None:
                break
            continue
        buf += line
        m = re.search(r"Step:\s*(\d+)", line)
        if m and int(m.group(1)) >= n:
            return buf
    raise AssertionError(f"step {n} never logged:\n{buf[-3000:]}")


def test_sigterm_saves_and_exits_cleanly_single_process(tmp_path):
    proc = _start(["--total_train_steps", "100000", "--work_dir", str(tmp_path)])
    try:
        out = _wait_for_step(proc, 3)
        proc.send_signal(signal.SIGTERM)
        rest, _ = proc.communicate(timeout=120)
    finally:
        if proc.poll() is None:
            os.killpg(proc.pid, signal.SIGKILL)
    text = out + rest
    assert proc.returncode == 0, text[-3000:]
    m = re.search(r"stop requested by signal: checkpoint written at step (\d+)", text)
    assert m, text[-3000:]
    step = int(m.group(1))
    assert 3 <= step < 1000
    files = os.listdir(tmp_path / str(step))
    assert any(f.startswith("weights_") for f in files) and "scheduler.pt" in files and not any(f.endswith(".tmp") for f in files)
    # and the job continues from there
    proc2 = _start(["--total_train_steps", str(step + 2), "--work_dir", str(tmp_path), "--resume_path", str(tmp_path / str(step))])
    out2, _ = proc2.communicate(timeout=300)
    assert proc2.returncode == 0, out2[-3000:]
    assert [int(s) for s in re.findall(r"Step:\s*(\d+)", out2)] == [step + 1, step + 2]


def test_sigterm_two_ranks_stop_after_the_same_step(tmp_path):
    proc = _start(["--total_train_steps", "100000", "--work_dir", str(tmp_path), "--data_parallel_size", "2",
                   "--signal_check_interval", "2"], nproc=2)
    try:
        out = _wait_for_step(proc, 3)
        os.killpg(proc.pid, signal.SIGTERM)                       # the launcher and both workers, as a scheduler would
        rest, _ = proc.communicate(timeout=180)
    finally:
        if proc.poll() is None:
            os.killpg(proc.pid, signal.SIGKILL)
    text = out + rest
    m = re.search(r"stop requested by signal: checkpoint written at step (\d+)", text)
    assert m, text[-3000:]
    step = int(m.group(1))
    assert step % 2 == 0                                          # agreement happens on the check interval
    assert any(f.startswith("weights_") for f in os.listdir(tmp_path / str(step)))
    assert "Traceback" not in text.split("stop requested by signal")[0]
