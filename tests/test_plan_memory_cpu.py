"""``tools/plan_memory.py``: the static part is exact (checked against real sharded models in
``test_parallel_cpu.py::test_replicated_parameters_start_equal``); here: totals for the vendored configs, the direction of
every knob, and the command line."""
import importlib.util
import json
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("plan_memory", os.path.join(ROOT, "tools", "plan_memory.py"))
pm = importlib.util.module_from_spec(spec)
spec.loader.exec_module(pm)


def _bytes(p, key, stage=0):
    return p["stages"][stage]["bytes"][key]


def test_parameter_counts_of_the_vendored_models():
    # totals quoted on the model cards (tied embeddings counted once)
    for name, billions in (("Qwen3-0.6B", 0.596), ("Qwen3-1.7B", 1.72), ("Qwen3-8B", 8.19), ("Qwen3-14B", 14.77),
                           ("Qwen3-32B", 32.76), ("Qwen3-30B-A3B", 30.53)):
        p = pm.plan(name)
        assert abs(p["stages"][0]["params"] / 1e9 - billions) < 0.01 * billions + 0.005, (name, p["stages"][0]["params"])
        b = p["stages"][0]["bytes"]
        assert b["params_bf16"] == 2 * p["stages"][0]["params"] and b["grads_fp32"] == 4 * p["stages"][0]["params"]
        assert b["adamw_state"] == 8 * p["stages"][0]["params"] and p["optimizer"] == "replicated AdamW"


def test_headline_and_target_layouts_fit_and_one_gpu_32b_does_not():
    assert pm.plan("Qwen3-0.6B", micro_batch_size=2, sequence_length=2048)["fits"]
    assert pm.plan("Qwen3-8B", tp=2, cp=2, dp=2, sp=True, sequence_length=4096)["fits"]
    # dropless EP8 sizes every row-padded activation for the worst case (ep x T x k rows) in all 48 layers: it does not fit,
    # which is what happened on the box -- the EP8 run in profiles/ used capacity factor 2.0, and that fits
    assert not pm.plan("Qwen3-30B-A3B", ep=8, sequence_length=4096)["fits"]
    assert pm.plan("Qwen3-30B-A3B", ep=8, sequence_length=4096, ep_capacity_factor=2.0)["fits"]
    assert pm.plan("Qwen3-30B-A3B", ep=8, sequence_length=4096, gradient_checkpointing=True)["fits"]
    assert pm.plan("Qwen3-32B", tp=4, pp=2, ga=8, sp=True, sequence_length=4096)["fits"]
    big = pm.plan("Qwen3-32B", sequence_length=4096)
    assert not big["fits"] and big["stages"][0]["fraction_of_hbm"] > 2.0


def test_every_knob_moves_memory_in_the_right_direction():
    base = pm.plan("Qwen3-8B", sequence_length=4096)
    tp2 = pm.plan("Qwen3-8B", tp=2, sequence_length=4096)
    assert abs(_bytes(tp2, "params_bf16") / _bytes(base, "params_bf16") - 0.5) < 0.01        # norms are replicated: just above half
    assert _bytes(tp2, "activations_est") < _bytes(base, "activations_est") and _bytes(tp2, "symmetric_windows") > 0
    sp = pm.plan("Qwen3-8B", tp=2, sp=True, sequence_length=4096)
    assert _bytes(sp, "activations_est") < _bytes(tp2, "activations_est")
    gc = pm.plan("Qwen3-8B", sequence_length=4096, gradient_checkpointing=True)
    assert _bytes(gc, "activations_est") < 0.2 * _bytes(base, "activations_est")
    cp2 = pm.plan("Qwen3-8B", cp=2, sequence_length=4096)
    assert abs(_bytes(cp2, "activations_est") / _bytes(base, "activations_est") - 0.5) < 0.01 and _bytes(cp2, "logits") * 2 == _bytes(base, "logits")
    dp8 = pm.plan("Qwen3-8B", dp=8, sequence_length=4096)
    assert dp8["optimizer"].startswith("zero1 on peer memory") and _bytes(dp8, "adamw_state") * 8 == _bytes(base, "adamw_state")
    assert pm.plan("Qwen3-8B", dp=8, zero_stage=0)["stages"][0]["bytes"]["adamw_state"] == _bytes(base, "adamw_state")
    long = pm.plan("Qwen3-8B", sequence_length=8192)
    assert _bytes(long, "activations_est") == 2 * _bytes(base, "activations_est")
    # pipeline: parameters split over the stages, stage 0 holds the most micro-batches (1F1B)
    pp2 = pm.plan("Qwen3-8B", pp=2, ga=4, sequence_length=4096)
    assert len(pp2["stages"]) == 2 and pp2["stages"][0]["micro_batches_in_flight"] == 2 and pp2["stages"][1]["micro_batches_in_flight"] == 1
    assert sum(st["params"] for st in pp2["stages"]) == base["stages"][0]["params"]          # untied: nothing is duplicated
    assert _bytes(pp2, "logits", 0) == 0 and _bytes(pp2, "logits", 1) == _bytes(base, "logits")
    tied = pm.plan("Qwen3-0.6B", pp=2)
    assert sum(st["params"] for st in tied["stages"]) == pm.plan("Qwen3-0.6B")["stages"][0]["params"] + 151936 * 1024   # tied head: a copy per end
    # experts: sharded over EP, their AdamW state is not sharded by the peer-memory optimiser
    ep8 = pm.plan("Qwen3-30B-A3B", ep=8)
    one = pm.plan("Qwen3-30B-A3B")
    assert abs(ep8["stages"][0]["params_expert"] * 8 - one["stages"][0]["params_expert"]) == 0
    assert _bytes(ep8, "adamw_state") > 8 * ep8["stages"][0]["params_expert"]


def test_invalid_layouts_are_refused_and_the_cli_prints_a_table(capsys):
    for kw in ({"tp": 3}, {"cp": 3, "sequence_length": 4096}, {"ep": 2}):
        with pytest.raises(SystemExit):
            pm.plan("Qwen3-8B", **kw)
    assert pm.main(["--model", "Qwen3-0.6B", "--micro_batch_size", "2", "--sequence_length", "2048"]) == 0
    out = capsys.readouterr().out
    assert "FITS in 180 GB" in out and "act(est)" in out
    assert pm.main(["--model", "Qwen3-32B", "--json"]) == 1
    d = json.loads(capsys.readouterr().out)
    assert d["fits"] is False and d["layout"]["world"] == 1 and d["stages"][0]["layers"] == 64


def test_suggest_lists_only_legal_layouts_that_fit():
    found = pm.suggest("Qwen3-32B", 8, sequence_length=4096)
    assert found and all(f["tp"] * f["pp"] * f["cp"] * f["ep"] * f["dp"] == 8 for f in found)
    assert all(f["ep"] == 1 for f in found)                                  # a dense model has no expert axis
    assert not any(f["dp"] == 8 for f in found)                              # 32B does not fit one GPU's 180 GB as pure DP
    assert all(f["worst_fraction_of_hbm"] <= 1.0 for f in found)
    assert found[0]["combination_covered_by_tests"] and found[0]["dp"] == max(f["dp"] for f in found)
    assert {(f["tp"], f["pp"]) for f in found} >= {(4, 2), (8, 1), (2, 4)}
    small = pm.suggest("Qwen3-0.6B", 8, micro_batch_size=2, sequence_length=2048)
    assert small[0] == {**small[0], "dp": 8, "tp": 1, "pp": 1, "cp": 1, "ep": 1}
    moe = pm.suggest("Qwen3-30B-A3B", 8, sequence_length=4096, ep_capacity_factor=2.0)
    assert any(f["ep"] == 8 for f in moe) and all(128 % f["ep"] == 0 and f["tp"] <= 4 for f in moe)   # 4 kv heads bound tp
    assert pm.main(["--model", "Qwen3-8B", "--suggest", "--gpus", "2", "--json"]) == 0
