"""CPU/gloo smoke test of the FSDP2 examples: `python -m pytest examples/FSDP2/test_smoke.py -q`."""
import os
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


def _run(script, *args, nproc=2):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1", "--nproc-per-node", str(nproc),
           os.path.join(HERE, script), *args]
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    p = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=600)
    assert p.returncode == 0, p.stderr[-3000:]
    return p.stdout


def test_fsdp2_trains():
    out = _run("fsdp2_main.py", "--steps", "3", "--explicit-prefetch", "--ac")
    assert "step   2 loss" in out


def test_fsdp2_checkpoint_resume(tmp_path):
    _run("fsdp2_main.py", "--steps", "2", "--ckpt-dir", str(tmp_path))
    out = _run("fsdp2_main.py", "--steps", "3", "--ckpt-dir", str(tmp_path))
    assert "resumed from step 2" in out and "step   2 loss" in out


@pytest.mark.parametrize("nproc,tp", [(2, 2), (4, 2)])
def test_fsdp2_tp(nproc, tp):
    """(4, 2) is the 2-D case of the reference's test_smoke_mp.py: dp2 x tp2 on four CPU processes (gloo)."""
    out = _run("fsdp2_tp_main.py", "--tp", str(tp), "--steps", "2", nproc=nproc)
    assert "step   1 loss" in out
