"""End-to-end runs of the trainer CLI (``tools/train.py``) on CPU / gloo: single process, 2-process layouts,
checkpoint save → resume, profiler window, and the watchdog with an injected hang."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LINE = re.compile(r"Step:\s*(\d+).*?Loss:\s*([\d.]+)")


def _run(args, nproc=1, env=None, timeout=600, expect_rc=0):
    base = [sys.executable]
    if nproc > 1:
        base += ["-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1", "--nproc-per-node", str(nproc)]
    cmd = base + [os.path.join(ROOT, "tools", "train.py"), "--synthetic_data", "true", "--use_cpu", "true",
                  "--sequence_length", "64", "--micro_batch_size", "2", "--learning_rate", "1e-3"] + args
    e = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2")
    e.update(env or {})
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=e, cwd=ROOT)
    assert p.returncode == expect_rc, f"rc={p.returncode}\n{p.stdout[-2000:]}\n{p.stderr[-3000:]}"
    return p.stdout + p.stderr


def _run_many(jobs):
    """Runs several independent CLI jobs ``(args, nproc)`` concurrently (process start-up dominates their cost)."""
    from concurrent.futures import ThreadPoolExecutor

    with ThreadPoolExecutor(max_workers=len(jobs)) as ex:
        futs = [ex.submit(_run, a, n) for a, n in jobs]
        return [f.result() for f in futs]


def _losses(out):
    return [(int(s), float(l)) for s, l in LINE.findall(out)]


def test_single_process_trains(tmp_path):
    out = _run(["--model_name_or_path", "tiny-llama", "--total_train_steps", "6", "--work_dir", str(tmp_path),
                "--profile_start_step", "2", "--profile_end_step", "4", "--trace_dir", str(tmp_path / "tr")])
    ls = _losses(out)
    assert [s for s, _ in ls] == [1, 2, 3, 4, 5, 6], out[-1500:]
    assert all(0.0 < l < 12.0 for _, l in ls)          # random tokens: the loss stays near ln(vocab)
    assert (tmp_path / "tr" / "trace_rank0.json").exists()
    assert any(f.startswith("performance_logs_0_") for f in os.listdir(tmp_path))


def test_two_process_layouts(tmp_path):
    layouts = {"tp2sp": ["--tensor_parallel_size", "2", "--sequence_parallel", "true"],
               "dp2ga2": ["--data_parallel_size", "2", "--gradient_accumulation_steps", "2"],
               "pp2ga2": ["--pipeline_parallel_size", "2", "--gradient_accumulation_steps", "2"],
               "cp2": ["--context_parallel_size", "2"],
               "dp2zero1": ["--data_parallel_size", "2", "--zero_stage", "1", "--save_model_checkpoint", "true",
                            "--save_frequency", "2"]}
    jobs = [(["--model_name_or_path", "tiny-qwen3", "--num_attention_heads", "4", "--num_key_value_heads", "2",
              "--total_train_steps", "4", "--work_dir", str(tmp_path / k)] + v, 2) for k, v in layouts.items()]
    for k, out in zip(layouts, _run_many(jobs)):
        ls = _losses(out)
        assert len(ls) == 4 and all(0.0 < l < 12.0 for _, l in ls), (k, out[-2000:])


def test_native_loader_cli(tmp_path):
    from scaletorch_b200.data import native_loader_available

    if not native_loader_available():
        pytest.skip("native library not built")
    out = _run(["--model_name_or_path", "tiny-llama", "--total_train_steps", "4", "--work_dir", str(tmp_path),
                "--native_loader", "true", "--num_samples", "64", "--data_parallel_size", "2",
                "--gradient_accumulation_steps", "2"], nproc=2)
    ls = _losses(out)
    assert len(ls) == 4 and all(0.0 < l < 12.0 for _, l in ls), out[-2000:]


def test_checkpoint_resume(tmp_path):
    const_lr = ["--lr_scheduler_type", "step", "--step_size", "1000"]     # the schedule must not depend on total steps
    common = ["--model_name_or_path", "tiny-llama", "--work_dir", str(tmp_path), "--save_model_checkpoint", "true",
              "--save_frequency", "3", "--log_lag", "0"] + const_lr
    out1, out3 = _run_many([(common + ["--total_train_steps", "3"], 1),
                            (["--model_name_or_path", "tiny-llama", "--work_dir", str(tmp_path / "straight"), "--log_lag", "0",
                              "--total_train_steps", "5"] + const_lr, 1)])
    assert (tmp_path / "3").is_dir()
    out2 = _run(common + ["--total_train_steps", "5", "--resume_path", str(tmp_path / "3")])
    assert "Loaded checkpoint at step 3" in out2
    assert [s for s, _ in _losses(out2)] == [4, 5]
    assert all(0.0 < l < 12.0 for _, l in _losses(out1) + _losses(out2))
    # a resumed run continues the uninterrupted trajectory: weights, optimiser state, LR schedule AND data position
    straight = dict(_losses(out3))
    for s, l in _losses(out2):
        assert abs(l - straight[s]) < 2e-3, (s, l, straight[s])


def test_watchdog_turns_a_hang_into_an_exit(tmp_path):
    out = _run(["--model_name_or_path", "tiny-llama", "--total_train_steps", "50", "--work_dir", str(tmp_path),
                "--step_timeout_s", "3"], env={"SCALETORCH_B200_FAULT_INJECT": "0:3:hang"}, expect_rc=124, timeout=120)
    assert "[watchdog] rank 0 stuck after step 3" in out
    assert (tmp_path / "hang_rank0.txt").exists() and (tmp_path / "heartbeat_rank0.json").exists()


def test_cli_loss_trajectories_agree_across_layouts(tmp_path):
    """Same seed, same global batch: single process, DP2 and (zig-zag) CP2 must print the same losses step by step."""
    common = ["--model_name_or_path", "tiny-qwen3", "--total_train_steps", "4", "--log_lag", "0", "--seed", "7",
              "--num_samples", "64"]
    outs = _run_many([(common + ["--micro_batch_size", "4", "--work_dir", str(tmp_path / "a")], 1),
                      (common + ["--micro_batch_size", "2", "--data_parallel_size", "2", "--work_dir", str(tmp_path / "b")], 2),
                      (common + ["--micro_batch_size", "4", "--context_parallel_size", "2", "--work_dir", str(tmp_path / "c")], 2)])
    single, dp2, cp2 = (_losses(o) for o in outs)
    assert len(single) == len(dp2) == len(cp2) == 4
    for (s, a), (_, b), (_, c) in zip(single, dp2, cp2):
        assert abs(a - b) < 2e-3 and abs(a - c) < 2e-3, (s, a, b, c)      # the metrics line prints 4 decimals


def test_checkpoint_resume_under_tensor_parallel(tmp_path):
    """Per-(tp, pp, ep) checkpoint files: a TP2+SP run resumed from step 3 continues the uninterrupted trajectory."""
    base = ["--model_name_or_path", "tiny-qwen3", "--num_attention_heads", "4", "--num_key_value_heads", "2",
            "--tensor_parallel_size", "2", "--sequence_parallel", "true", "--log_lag", "0",
            "--lr_scheduler_type", "step", "--step_size", "1000"]
    saving = ["--save_model_checkpoint", "true", "--save_frequency", "3", "--work_dir", str(tmp_path / "run")]
    _, straight_out = _run_many([(base + saving + ["--total_train_steps", "3"], 2),
                                 (base + ["--total_train_steps", "5", "--work_dir", str(tmp_path / "straight")], 2)])
    files = sorted(os.listdir(tmp_path / "run" / "3"))
    assert any(f.startswith("weights_tp_rank_world_size=0_2") for f in files) and any(
        f.startswith("weights_tp_rank_world_size=1_2") for f in files), files
    resumed = _losses(_run(base + saving + ["--total_train_steps", "5", "--resume_path", str(tmp_path / "run" / "3")], nproc=2))
    straight = dict(_losses(straight_out))
    assert [s for s, _ in resumed] == [4, 5]
    for s, l in resumed:
        assert abs(l - straight[s]) < 2e-3, (s, l, straight[s])


def test_async_checkpoint_cli(tmp_path):
    """--async_checkpoint true: every checkpoint is published (no temporary file left) by the time the job exits, and a
    run resumed from the last one continues."""
    common = ["--model_name_or_path", "tiny-llama", "--work_dir", str(tmp_path), "--save_model_checkpoint", "true",
              "--save_frequency", "2", "--async_checkpoint", "true", "--log_lag", "0", "--lr_scheduler_type", "step", "--step_size", "1000"]
    out = _run(common + ["--total_train_steps", "4"])
    assert len(_losses(out)) == 4
    for step in ("2", "4"):
        files = os.listdir(tmp_path / step)
        assert any(f.startswith("weights_") and f.endswith(".pth") for f in files) and not any(f.endswith(".tmp") for f in files)
    out2 = _run(common + ["--total_train_steps", "5", "--resume_path", str(tmp_path / "4")])
    assert [s for s, _ in _losses(out2)] == [5]


def test_periodic_evaluation_cli(tmp_path):
    """``--eval_interval``: evaluation lines appear at the right steps, every token of the held-out set is
    counted once under TP2+SP and CP2, the held-out loss under CP2 (same initial weights, same global batch) equals the
    single-process one, evaluation leaves the training
    trajectory untouched, and a token file works as the held-out set."""
    import numpy as np

    ev = re.compile(r"Eval \| Step:\s*(\d+)\s*\| Loss:\s*([\d.]+).*?Tokens:\s*(\d+)")
    common = ["--model_name_or_path", "tiny-qwen3", "--num_attention_heads", "4", "--num_key_value_heads", "2",
              "--total_train_steps", "4", "--save_frequency", "0"]
    evalf = ["--eval_interval", "2", "--eval_batches", "2"]
    tok = tmp_path / "heldout.npy"
    np.save(tok, np.random.default_rng(0).integers(0, 2048, size=65 * 9, dtype=np.int32))
    jobs = [(common + ["--work_dir", str(tmp_path / "plain")], 1),
            (common + evalf + ["--work_dir", str(tmp_path / "single")], 1),
            (common + evalf + ["--work_dir", str(tmp_path / "tp2"), "--tensor_parallel_size", "2", "--sequence_parallel", "true"], 2),
            (common + evalf + ["--work_dir", str(tmp_path / "cp2"), "--context_parallel_size", "2"], 2),
            (common + evalf + ["--work_dir", str(tmp_path / "file"), "--eval_token_file", str(tok)], 1),
            (common + evalf + ["--work_dir", str(tmp_path / "dp2"), "--data_parallel_size", "2", "--micro_batch_size", "1"], 2)]
    plain, single, tp2, cp2, from_file, dp2 = _run_many(jobs)

    def train_losses(out):
        return _losses("\n".join(l for l in out.splitlines() if "Eval |" not in l))

    assert not ev.findall(plain)
    assert train_losses(single) == train_losses(plain) and len(train_losses(plain)) == 4    # evaluation does not perturb training
    e1 = [(int(s), float(l), int(t)) for s, l, t in ev.findall(single)]
    assert [s for s, _, _ in e1] == [2, 4] and all(t == 2 * 2 * 64 for _, _, t in e1), single[-1500:]
    for name, out in (("tp2", tp2), ("cp2", cp2), ("dp2", dp2)):
        e = [(int(s), float(l), int(t)) for s, l, t in ev.findall(out)]
        assert [(s, t) for s, _, t in e] == [(s, t) for s, _, t in e1], (name, out[-1500:])
        if name in ("cp2", "dp2"):                                 # same initial weights, same global batch: same model
            assert all(abs(a[1] - b[1]) < 2e-3 for a, b in zip(e, e1)), (name, e, e1)
        else:                                                      # tensor-parallel shards are seeded per rank: another model
            assert all(0.0 < l < 12.0 for _, l, _ in e), (name, e)
    ef = [(int(s), float(l), int(t)) for s, l, t in ev.findall(from_file)]
    assert [s for s, _, _ in ef] == [2, 4] and ef[0][2] == 2 * 2 * 64 and all(0 < l < 12 for _, l, _ in ef)
    assert train_losses(from_file) == train_losses(plain)


def test_evaluation_flags_are_validated():
    from scaletorch_b200.trainer import ScaleTorchArguments

    with pytest.raises(ValueError, match="eval_interval"):
        ScaleTorchArguments(model_name_or_path="tiny-llama", eval_interval=-1)
    with pytest.raises(ValueError, match="eval_batches"):
        ScaleTorchArguments(model_name_or_path="tiny-llama", eval_batches=0)


def test_checkpoint_retention_and_resume_latest(tmp_path):
    """``--keep_last_checkpoints 2`` leaves the two newest checkpoints; ``--resume_path latest`` continues from the newest one
    (and starts from scratch in an empty work directory) with the same losses as an uninterrupted run."""
    common = ["--model_name_or_path", "tiny-llama", "--save_frequency", "2", "--log_lag", "0",
              "--lr_scheduler_type", "step", "--step_size", "1000"]       # the schedule must not depend on total steps
    full = _losses(_run(common + ["--total_train_steps", "10", "--work_dir", str(tmp_path / "full"), "--save_frequency", "0"]))
    wd = tmp_path / "wd"
    first = _run(common + ["--total_train_steps", "8", "--work_dir", str(wd), "--keep_last_checkpoints", "2",
                           "--resume_path", "latest"])          # nothing to resume from yet: starts at step 1
    assert [s for s, _ in _losses(first)] == list(range(1, 9))
    assert sorted(d for d in os.listdir(wd) if d.isdigit()) == ["6", "8"], os.listdir(wd)
    second = _run(common + ["--total_train_steps", "10", "--work_dir", str(wd), "--keep_last_checkpoints", "2",
                            "--resume_path", "latest"])
    assert "Loaded checkpoint at step 8" in second
    got = _losses(second)
    assert [s for s, _ in got] == [9, 10]
    assert all(abs(a[1] - b[1]) < 2e-3 for a, b in zip(got, full[8:])), (got, full[8:])
    assert sorted(d for d in os.listdir(wd) if d.isdigit()) == ["10", "8"]


def test_nonfinite_loss_aborts_before_a_checkpoint_is_written(tmp_path):
    """An exploding run stops with a non-zero exit code and a message that names the step; no checkpoint of the broken
    weights is published; ``--nonfinite_loss warn`` keeps the reference's behaviour (log and continue)."""
    boom = ["--model_name_or_path", "tiny-llama", "--learning_rate", "1e30", "--max_grad_norm", "0", "--total_train_steps", "6",
            "--save_frequency", "2"]
    out = _run(boom + ["--work_dir", str(tmp_path / "abort")], expect_rc=1)
    m = re.search(r"non-finite loss \((nan|inf)\) at step (\d+)", out)
    assert m and int(m.group(2)) == 2, out[-1500:]
    assert not any(d.isdigit() for d in os.listdir(tmp_path / "abort")), os.listdir(tmp_path / "abort")
    out = _run(boom + ["--work_dir", str(tmp_path / "warn"), "--nonfinite_loss", "warn", "--save_frequency", "0"])
    assert "WARNING: non-finite loss" in out and len(_losses(out.replace("nan", "0.0"))) >= 1
    from scaletorch_b200.trainer import ScaleTorchArguments

    with pytest.raises(ValueError, match="nonfinite_loss"):
        ScaleTorchArguments(model_name_or_path="tiny-llama", nonfinite_loss="ignore")


def test_resume_across_layouts_with_resharding(tmp_path):
    """A checkpoint written under TP2+SP, and one written under EP2 (MoE), are resumed in ONE process; a single-process
    checkpoint under PP2: refused without ``--resume_reshard true``; with it, the resumed steps have the losses of the
    uninterrupted run: weights, data position AND the AdamW moments (re-cut by name through the slot table stored in the
    checkpoint; expert moments included) carry over."""
    const = ["--log_lag", "0", "--lr_scheduler_type", "step", "--step_size", "1000"]
    dense = ["--model_name_or_path", "tiny-qwen3", "--num_attention_heads", "4", "--num_key_value_heads", "2"] + const
    moe = ["--model_name_or_path", "tiny-qwen3-moe"] + const
    tp2 = ["--tensor_parallel_size", "2", "--sequence_parallel", "true"]
    ep2 = ["--expert_parallel_size", "2", "--micro_batch_size", "1"]          # two data ranks x 1 = the single process's batch of 2
    _, straight, _, straight_moe = _run_many([
        (dense + tp2 + ["--total_train_steps", "3", "--save_frequency", "3", "--work_dir", str(tmp_path / "tp2")], 2),
        (dense + tp2 + ["--total_train_steps", "5", "--save_frequency", "0", "--work_dir", str(tmp_path / "straight")], 2),
        (moe + ep2 + ["--total_train_steps", "3", "--save_frequency", "3", "--work_dir", str(tmp_path / "ep2")], 2),
        (moe + ep2 + ["--total_train_steps", "5", "--save_frequency", "0", "--work_dir", str(tmp_path / "straight_moe")], 2)])
    want, want_moe = dict(_losses(straight)), dict(_losses(straight_moe))
    resume = ["--total_train_steps", "5", "--save_frequency", "5"]
    jobs = [(dense + ["--total_train_steps", "4", "--save_frequency", "0", "--work_dir", str(tmp_path / "refused"),
                      "--resume_path", str(tmp_path / "tp2" / "3")], 1),
            (dense + resume + ["--work_dir", str(tmp_path / "one"), "--resume_path", str(tmp_path / "tp2" / "3"), "--resume_reshard", "true"], 1),
            (moe + resume + ["--work_dir", str(tmp_path / "one_moe"), "--resume_path", str(tmp_path / "ep2" / "3"), "--resume_reshard", "true"], 1)]
    from concurrent.futures import ThreadPoolExecutor

    with ThreadPoolExecutor(max_workers=3) as ex:
        futs = [ex.submit(_run, a, n, None, 600, rc) for (a, n), rc in zip(jobs, (1, 0, 0))]
        refused, out, out_moe = [f.result() for f in futs]
    assert "--resume_reshard true" in refused
    assert "resharded the weights and AdamW moments" in out and "{'tp': 2, 'pp': 1, 'ep': 1}" in out and "Loaded checkpoint at step 3" in out
    got = dict(_losses(out))
    # step 4 depends on the weights and the data position, step 5 also on the re-cut AdamW moments and the step count
    assert sorted(got) == [4, 5] and abs(got[4] - want[4]) < 2e-3 and abs(got[5] - want[5]) < 2e-3, (got, want)
    got_moe = dict(_losses(out_moe))
    assert "{'tp': 1, 'pp': 1, 'ep': 2}" in out_moe and sorted(got_moe) == [4, 5]
    assert all(abs(got_moe[s] - want_moe[s]) < 2e-3 for s in (4, 5)), (got_moe, want_moe)
    # and onwards: the single-process checkpoint of step 5 continues under pipeline parallelism
    out = _run(dense + ["--total_train_steps", "6", "--save_frequency", "0", "--work_dir", str(tmp_path / "pp2"),
                        "--pipeline_parallel_size", "2", "--gradient_accumulation_steps", "1",
                        "--resume_path", str(tmp_path / "one" / "5"), "--resume_reshard", "true"], nproc=2)
    assert "resharded the weights" in out and [s for s, _ in _losses(out)] == [6]
    assert abs(dict(_losses(out))[6] - got[5]) < 1.0          # same model a step later: the loss stays where it was


def test_periodic_evaluation_through_pipeline_stages(tmp_path):
    """``--eval_interval`` with pipeline parallelism: the forward-only pass through the stages gives the held-out loss of the
    same weights evaluated in one process (both runs resume the same single-process checkpoint; the PP2 run re-cuts it)."""
    ev = re.compile(r"Eval \| Step:\s*(\d+)\s*\| Loss:\s*([\d.]+).*?Tokens:\s*(\d+)")
    model = ["--model_name_or_path", "tiny-qwen3", "--num_attention_heads", "4", "--num_key_value_heads", "2", "--log_lag", "0",
             "--lr_scheduler_type", "step", "--step_size", "1000"]
    _run(model + ["--total_train_steps", "2", "--save_frequency", "2", "--work_dir", str(tmp_path / "base")])
    resume = ["--total_train_steps", "4", "--save_frequency", "0", "--eval_interval", "2", "--eval_batches", "2",
              "--resume_path", str(tmp_path / "base" / "2")]
    one, pp2 = _run_many([(model + resume + ["--work_dir", str(tmp_path / "one")], 1),
                          (model + resume + ["--work_dir", str(tmp_path / "pp2"), "--pipeline_parallel_size", "2",
                                             "--resume_reshard", "true"], 2)])
    e1 = [(int(s), float(l), int(t)) for s, l, t in ev.findall(one)]
    e2 = [(int(s), float(l), int(t)) for s, l, t in ev.findall(pp2)]
    assert [(s, t) for s, _, t in e1] == [(4, 2 * 2 * 64)] == [(s, t) for s, _, t in e2], (one[-800:], pp2[-800:])
    assert abs(e1[0][1] - e2[0][1]) < 2e-3, (e1, e2)
