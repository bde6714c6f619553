"""``tools/supervise.py``: a training job killed mid-run is relaunched from the newest COMPLETE checkpoint and finishes
(reference: recovery is a manual relaunch with ``--resume_path``)."""
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _tool():
    spec = importlib.util.spec_from_file_location("supervise", os.path.join(ROOT, "tools", "supervise.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_complete_checkpoints_skips_partial_directories(tmp_path):
    sup = _tool()
    assert sup.complete_checkpoints(str(tmp_path / "missing")) == []
    for step, files in {2: ["weights_a.pth", "weights_b.pth"], 4: ["weights_a.pth", "weights_b.pth", "scheduler.pt"],
                        6: ["weights_a.pth"],                                  # a shard is missing: crashed while saving
                        8: ["weights_a.pth", "weights_b.pth", "weights_b.pth.tmp"],   # an unpublished temporary file
                        }.items():
        d = tmp_path / str(step)
        d.mkdir()
        for f in files:
            (d / f).write_text("x")
    (tmp_path / "notes").mkdir()
    got = sup.complete_checkpoints(str(tmp_path))
    assert [s for s, _ in got] == [2, 4] and got[-1][1].endswith("/4")


def test_with_resume_replaces_an_existing_flag():
    sup = _tool()
    cmd = ["python", "train.py", "--resume_path", "old", "--x", "1", "--resume_path=older"]
    assert sup.with_resume(cmd, "new") == ["python", "train.py", "--x", "1", "--resume_path", "new"]
    assert sup.with_resume(cmd, None) == ["python", "train.py", "--x", "1"]


def test_killed_job_is_resumed_and_finishes(tmp_path, monkeypatch, capfd):
    sup = _tool()
    work = tmp_path / "work"
    train = [sys.executable, os.path.join(ROOT, "tools", "train.py"), "--synthetic_data", "true", "--use_cpu", "true",
             "--model_name_or_path", "tiny-llama", "--sequence_length", "64", "--micro_batch_size", "2", "--learning_rate", "1e-3",
             "--lr_scheduler_type", "step", "--step_size", "1000", "--log_lag", "0", "--total_train_steps", "6",
             "--work_dir", str(work), "--save_model_checkpoint", "true", "--save_frequency", "2"]
    monkeypatch.setenv("CUDA_VISIBLE_DEVICES", "")
    monkeypatch.setenv("SCALETORCH_B200_FAULT_INJECT", "0:3:exit")            # rank 0 dies (exit code 17) before its 4th step
    rc = sup.main(["--work_dir", str(work), "--max_restarts", "2", "--grace_s", "2", "--"] + train)
    out = capfd.readouterr()
    text = out.out + out.err
    assert rc == 0, text[-3000:]
    attempts = [json.loads(l) for l in open(work / "supervisor.jsonl")]
    assert [a["rc"] for a in attempts] == [17, 0] and attempts[1]["resumed_from"] == 2
    assert "restart 1/2 from step 2" in text and "Loaded checkpoint at step 2" in text
    import re

    steps = [int(s) for s in re.findall(r"Step:\s*(\d+)", text)]
    assert steps == [1, 2, 3, 3, 4, 5, 6]                                      # step 3 ran twice: once before the crash, once after the resume
    assert (work / "6").is_dir()


def test_gives_up_when_restarts_make_no_progress(tmp_path):
    sup = _tool()
    rc = sup.main(["--work_dir", str(tmp_path), "--max_restarts", "5", "--max_stalls", "2", "--grace_s", "1", "--",
                   sys.executable, "-c", "import sys; sys.exit(9)"])
    assert rc == 3                                                             # not 5 blind restarts
    attempts = [json.loads(l) for l in open(tmp_path / "supervisor.jsonl")]
    assert len(attempts) == 2 and all(a["rc"] == 9 for a in attempts)
