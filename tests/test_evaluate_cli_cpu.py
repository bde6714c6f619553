"""``tools/evaluate.py``: the held-out loss is the same number under every layout (single process, DP2, CP2, TP2), counts
every token exactly once, and honours ``--max_batches``."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _eval(args, nproc=1):
    base = [sys.executable]
    if nproc > 1:
        base += ["-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1", "--nproc-per-node", str(nproc)]
    cmd = base + [os.path.join(ROOT, "tools", "evaluate.py"), "--use_cpu", "true", "--dtype", "float32", "--sequence_length", "64",
                  "--micro_batch_size", "4"] + args
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT,
                       env=dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2"))
    assert p.returncode == 0, p.stdout[-1500:] + p.stderr[-3000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("{")][-1]
    return json.loads(line)


@pytest.fixture(scope="module")
def token_file(tmp_path_factory):
    path = tmp_path_factory.mktemp("val") / "tokens.npy"
    np.save(path, np.random.RandomState(0).randint(0, 2048, size=40 * 65 + 17).astype(np.int32))   # 40 windows + a ragged tail
    return str(path)


@pytest.fixture(scope="module")
def model_dir(tmp_path_factory):
    """A tiny Qwen3 with 4 query / 2 kv heads and saved weights: every layout must load the SAME model (random init
    differs between tensor-parallel shards by design)."""
    import shutil

    import torch
    from safetensors.torch import save_file

    from scaletorch_b200.trainer import ScaleTorchArguments, create_model
    from scaletorch_b200.utils.checkpoint import export_hf_state_dict

    d = tmp_path_factory.mktemp("model") / "tiny"
    shutil.copytree(os.path.join(ROOT, "configs", "tiny-qwen3"), d)
    c = json.load(open(d / "config.json"))
    c.update(num_attention_heads=4, num_key_value_heads=2)
    json.dump(c, open(d / "config.json", "w"))
    torch.manual_seed(0)
    m, mc = create_model(ScaleTorchArguments(model_name_or_path=str(d), micro_batch_size=1, sequence_length=16, dtype="float32",
                                             use_cpu=True), torch.float32, torch.device("cpu"))
    save_file(export_hf_state_dict(m, mc), str(d / "model.safetensors"), metadata={"format": "pt"})
    return str(d)


def test_same_loss_under_every_layout(token_file, model_dir):
    from concurrent.futures import ThreadPoolExecutor

    jobs = {"single": ([], 1), "dp2": (["--data_parallel_size", "2"], 2), "cp2": (["--context_parallel_size", "2"], 2),
            "tp2": (["--tensor_parallel_size", "2"], 2)}
    with ThreadPoolExecutor(max_workers=4) as ex:
        common = ["--model_name_or_path", model_dir, "--load_pretrained_weights", "true", "--token_file", token_file]
        futs = {k: ex.submit(_eval, common + a, n) for k, (a, n) in jobs.items()}
        res = {k: f.result() for k, f in futs.items()}
    base = res["single"]
    assert base["tokens"] == 40 * 64 and base["batches"] == 10 and 7.0 < base["loss"] < 9.5
    assert abs(base["perplexity"] - np.exp(base["loss"])) < 1e-6 * base["perplexity"]
    for k in ("dp2", "cp2", "tp2"):
        assert res[k]["tokens"] == base["tokens"], (k, res[k])
        assert abs(res[k]["loss"] - base["loss"]) < 2e-4, (k, res[k]["loss"], base["loss"])
    assert res["dp2"]["batches"] == 5                              # each data rank saw half of the windows


def test_max_batches_and_synthetic_data(token_file, model_dir):
    r = _eval(["--model_name_or_path", model_dir, "--token_file", token_file, "--max_batches", "3"])
    assert r["batches"] == 3 and r["tokens"] == 3 * 4 * 64
    r = _eval(["--model_name_or_path", "tiny-llama", "--synthetic_data", "true", "--num_samples", "16", "--max_batches", "2"])
    assert r["tokens"] == 2 * 4 * 64


def test_generate_cli_is_deterministic_and_layout_independent(model_dir):
    """tools/generate.py: greedy continuation of the saved model, identical in one process and under TP2."""
    def gen(extra, nproc=1):
        base = [sys.executable]
        if nproc > 1:
            base += ["-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1", "--nproc-per-node", str(nproc)]
        cmd = base + [os.path.join(ROOT, "tools", "generate.py"), "--model_name_or_path", model_dir, "--load_pretrained_weights", "true",
                      "--use_cpu", "true", "--dtype", "float32", "--prompt_ids", "5,17,42,7", "--max_new_tokens", "6"] + extra
        p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT,
                           env=dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2"))
        assert p.returncode == 0, p.stdout[-1500:] + p.stderr[-3000:]
        return json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][-1])

    a = gen([])
    assert a["prompt_ids"] == [5, 17, 42, 7] and len(a["new_ids"]) == 6
    assert gen(["--tensor_parallel_size", "2"], nproc=2)["new_ids"] == a["new_ids"]
