"""The benchmark / verification tools run end to end on the host (reference: ``tools/bench_single.py``,
``tools/profile_mfu.py``, ``tools/optimize_mfu.py``, ``tools/verify_qwen3.py``, ``tests/models/benchmark_moe.py``).

On a GPU box they time bf16 steps with CUDA events; ``--cpu`` (or ``--device cpu`` / ``--no_run``) runs the same loop in
fp32 so that the plumbing -- model creation from a named config, the fused-loss call, optimiser construction, the report
format other scripts parse -- is checked without a GPU.
"""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, timeout=600):
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""), CUDA_VISIBLE_DEVICES="")
    p = subprocess.run([sys.executable, *args], cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    return p.stdout


def test_bench_single_cpu_dry_run_prints_the_report():
    out = _run("tools/bench_single.py", "--cpu", "--model_path", "tiny-llama", "--batch_size", "1", "--seq_len", "32", "--steps", "2")
    assert re.search(r"Model params: \d+\.\d+M", out)
    ms = float(re.search(r"Step time: ([\d.]+)ms", out).group(1))
    tps = float(re.search(r"Tokens/s: (\d+)", out).group(1))
    assert ms > 0 and abs(tps - 32 / (ms / 1e3)) / tps < 0.05
    loss = float(re.search(r"Avg Loss: ([\d.]+)", out).group(1))
    assert 5.0 < loss < 10.0                                       # ln(2048) = 7.6 at random init


def test_bench_single_without_a_gpu_says_what_to_do():
    env = dict(os.environ, PYTHONPATH=ROOT, CUDA_VISIBLE_DEVICES="")
    p = subprocess.run([sys.executable, "tools/bench_single.py", "--model_path", "tiny-llama"], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=300)
    assert p.returncode != 0 and "--cpu" in (p.stdout + p.stderr)


def test_optimize_mfu_sweeps_both_checkpointing_settings():
    out = _run("tools/optimize_mfu.py", "--cpu", "--model", "tiny-qwen3", "--batch_sizes", "1", "--seq_lens", "32", "--steps", "1")
    rows = [l.split() for l in out.splitlines() if re.match(r"^1\s+32\s+(True|False)", l)]
    assert sorted(r[2] for r in rows) == ["False", "True"]
    assert all(float(r[3]) > 0 and float(r[4]) > 0 for r in rows)


def test_profile_mfu_analytic_breakdown_adds_up():
    out = _run("tools/profile_mfu.py", "--model", "Qwen3-0.6B", "--batch_size", "2", "--seq_len", "2048", "--no_run")
    val = {m.group(1): float(m.group(2)) for m in re.finditer(r"^(\w+)\s+([\d.]+) T", out, re.M)}
    assert abs(val["fwd_linear"] + val["fwd_attention"] + val["fwd_lm_head"] - val["fwd_total"]) < 2e-3
    assert abs(val["train_total"] - 3 * val["fwd_total"]) < 5e-3     # backward = 2 x forward


def test_verify_qwen3_on_the_builtin_tiny_config():
    out = _run("tools/verify_qwen3.py", "--cpu")
    assert "tied embedding/lm_head share storage: True" in out
    assert "forward finite: True" in out and "gradients finite and non-zero: True" in out


def test_benchmark_moe_compares_dense_and_moe():
    out = _run("tools/benchmark_moe.py", "--device", "cpu", "--n_embd", "32", "--n_layer", "2", "--n_head", "2", "--n_exp", "4",
               "--batch", "1", "--block", "16", "--iters", "1")
    assert re.search(r"^dense\s+[\d.]+", out, re.M) and re.search(r"^moe4x top2\s+[\d.]+", out, re.M)
    assert "'moe_layer_count': 1" in out
