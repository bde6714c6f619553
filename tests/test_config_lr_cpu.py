"""Every validator of the training dataclasses and the closed-form shape of every LR schedule (reference suites:
``tests/trainer/test_config.py``, ``tests/trainer/test_lr_scheduler.py``)."""
import dataclasses
import math
import os

import pytest
import torch

from scaletorch_b200.trainer import config as K
from scaletorch_b200.trainer.lr_scheduler import create_lr_scheduler, register_scheduler, registered_schedulers


# ------------------------------------------------------------------------------------------- argument groups
def test_group_defaults():
    d = K.DataArguments()
    assert (d.data_path, d.dataset_name, d.split, d.num_proc, d.num_workers, d.num_samples, d.pin_memory) == \
        ("./data", "wikitext2", "train", 1, 2, None, True) and d.synthetic_data is False
    p = K.ParallelArguments()
    assert (p.tensor_parallel_size, p.pipeline_parallel_size, p.data_parallel_size, p.context_parallel_size,
            p.expert_parallel_size) == (1, 1, 1, 1, 1)
    assert p.fused_comm and p.cp_zigzag and not p.sequence_parallel and p.ep_capacity_factor is None
    s = K.LrSchedulerArguments()
    assert (s.lr_scheduler_type, s.warmup_steps, s.T_max, s.eta_min, s.power, s.step_size, s.gamma, s.pct_start) == \
        ("linear", 0, None, 0.0, 1.0, 1, 0.1, 0.3)
    o = K.OptimizerArguments()
    assert o.optimizer_type == "adamw" and o.betas == (0.9, 0.999) and o.weight_decay == 0.0 and o.use_fused_adam
    t = K.TrainingArguments()
    assert (t.batch_size, t.gradient_accumulation_steps, t.epochs, t.seed, t.sequence_length, t.log_interval) == \
        (64, 1, 5, 1, 1024, 100) and t.micro_batch_size is None and t.max_grad_norm == 1.0 and t.zero_stage == 0
    c = K.CheckpointArguments()
    assert (c.work_dir, c.save_model_checkpoint, c.save_frequency, c.resume_path) == ("./work_dir", True, 300, "")
    g = K.LoggingArguments()
    assert not g.use_wandb and g.project_name == "scaletorch" and g.experiment_name is None


def test_model_arguments_fill_from_vendored_config_and_tolerate_unknown_models():
    m = K.ModelArguments(model_name_or_path="Qwen3-0.6B")
    assert (m.num_hidden_layers, m.num_attention_heads, m.num_key_value_heads) == (28, 16, 8)
    m = K.ModelArguments(model_name_or_path="Qwen3-0.6B", num_hidden_layers=4)     # explicit overrides win
    assert m.num_hidden_layers == 4 and m.num_attention_heads == 16
    m = K.ModelArguments(model_name_or_path="/no/such/model")
    assert m.num_hidden_layers is None and m.num_attention_heads is None
    assert K.resolve_model_config("/no/such/model", quiet=True) is None
    with pytest.raises(RuntimeError, match="cannot resolve"):
        K.resolve_model_config("/no/such/model")


@pytest.mark.parametrize("field", ["tensor_parallel_size", "pipeline_parallel_size", "data_parallel_size",
                                   "context_parallel_size", "expert_parallel_size"])
@pytest.mark.parametrize("bad", [0, -2])
def test_parallel_sizes_must_be_positive(field, bad):
    with pytest.raises(ValueError, match=field):
        K.ParallelArguments(**{field: bad})


def test_parallel_enums_and_env(monkeypatch):
    for kw in (dict(pipeline_parallel_engine="gpipe"), dict(backend="mpi"), dict(backend="hccl"), dict(grad_wire_dtype="fp8")):
        with pytest.raises(ValueError):
            K.ParallelArguments(**kw)
    assert K.ParallelArguments(pipeline_parallel_engine="afab", backend="gloo", grad_wire_dtype="bfloat16").backend == "gloo"
    monkeypatch.setenv("SEQUENCE_PARALLEL", "1")                  # the reference's env switch still works
    assert K.ParallelArguments().sequence_parallel
    monkeypatch.setenv("SEQUENCE_PARALLEL", "0")
    assert not K.ParallelArguments().sequence_parallel


@pytest.mark.parametrize("kw", [dict(lr_scheduler_type="exponential"), dict(warmup_steps=-1),
                                dict(lr_scheduler_type="cosine", eta_min=-0.1),
                                dict(lr_scheduler_type="polynomial", power=0.0), dict(lr_scheduler_type="polynomial", power=-1.0),
                                dict(lr_scheduler_type="step", step_size=0), dict(lr_scheduler_type="step", gamma=0.0),
                                dict(lr_scheduler_type="step", gamma=1.01), dict(lr_scheduler_type="onecycle", pct_start=0.0),
                                dict(lr_scheduler_type="onecycle", pct_start=1.0)])
def test_lr_argument_validators(kw):
    with pytest.raises(ValueError):
        K.LrSchedulerArguments(**kw)


def test_lr_validators_only_bind_for_their_own_scheduler():
    K.LrSchedulerArguments(lr_scheduler_type="linear", gamma=5.0, power=-1.0, pct_start=7.0)    # unused fields are not policed
    K.LrSchedulerArguments(lr_scheduler_type="step", gamma=1.0, step_size=3)


@pytest.mark.parametrize("kw", [dict(optimizer_type="adagrad"), dict(betas=[0.9]), dict(betas=[0.9, 0.95, 0.99])])
def test_optimizer_validators(kw):
    with pytest.raises(ValueError):
        K.OptimizerArguments(**kw)


def test_betas_become_a_float_tuple():
    assert K.OptimizerArguments(betas=["0.8", 0.9]).betas == (0.8, 0.9)


@pytest.mark.parametrize("cls,kw", [(K.TrainingArguments, dict(gradient_accumulation_steps=0)),
                                    (K.TrainingArguments, dict(micro_batch_size=0)),
                                    (K.TrainingArguments, dict(sequence_length=0)),
                                    (K.CheckpointArguments, dict(save_frequency=-1))])
def test_training_and_checkpoint_validators(cls, kw):
    with pytest.raises(ValueError):
        cls(**kw)


def test_combined_arguments_batch_arithmetic():
    c = K.ScaleTorchArguments()
    assert c.micro_batch_size == c.batch_size                     # micro batch falls back to batch_size
    c = K.ScaleTorchArguments(micro_batch_size=2, gradient_accumulation_steps=4, data_parallel_size=3, sequence_length=512)
    assert c.global_batch_size == 24 and c.global_batch_size_token == 24 * 512
    c = K.ScaleTorchArguments(micro_batch_size=2, data_parallel_size=2, expert_parallel_size=4)
    assert c.global_batch_size == 16                              # EP ranks read distinct samples by default ...
    c = K.ScaleTorchArguments(micro_batch_size=2, data_parallel_size=2, expert_parallel_size=4, ep_replicate_data=True)
    assert c.global_batch_size == 4                               # ... unless the reference's replicated reading is asked for
    c = K.ScaleTorchArguments(sequence_length=None)
    assert c.global_batch_size_token is None
    with pytest.raises(ValueError, match="divisible"):
        K.ScaleTorchArguments(sequence_length=1025, context_parallel_size=2)


def test_world_size_validation_message():
    c = K.ScaleTorchArguments(tensor_parallel_size=2, pipeline_parallel_size=2, data_parallel_size=2)
    c.validate_world_size(8)
    K.validate_config(c, 8)
    with pytest.raises(ValueError, match=r"world_size \(4\).*= 8"):
        c.validate_world_size(4)


def test_every_field_is_a_cli_flag_with_help():
    from transformers import HfArgumentParser

    parser = HfArgumentParser(K.ScaleTorchArguments)
    flags = {a.dest for a in parser._actions}
    for f in dataclasses.fields(K.ScaleTorchArguments):
        if not f.init:
            continue
        assert f.name in flags, f.name
        assert f.metadata.get("help"), f"--{f.name} has no help text"
    names = [f.name for f in dataclasses.fields(K.ScaleTorchArguments)]
    assert len(names) == len(set(names))                          # the eight groups do not shadow each other


# ------------------------------------------------------------------------------------------- schedules, closed form
def _run(kind, steps=40, lr=2.0, **kw):
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=lr)
    cfg = K.LrSchedulerArguments(lr_scheduler_type=kind, **kw)
    sched = create_lr_scheduler(opt, cfg, steps)
    out = []
    for _ in range(steps):
        out.append(opt.param_groups[0]["lr"])
        opt.step()
        sched.step()
    return out


def test_registry_lists_builtins():
    assert {"constant", "linear", "cosine", "polynomial", "step", "onecycle"} <= set(registered_schedulers())


def test_warmup_is_linear_from_one_step():
    lrs = _run("constant", warmup_steps=4)
    assert lrs[:5] == pytest.approx([0.5, 1.0, 1.5, 2.0, 2.0])


def test_linear_reaches_zero_at_the_end_and_stays_there():
    lrs = _run("linear", steps=20, warmup_steps=0)
    assert lrs[0] == 2.0 and lrs[10] == pytest.approx(1.0) and lrs[19] == pytest.approx(0.1)
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=1.0)
    sched = create_lr_scheduler(opt, K.LrSchedulerArguments(lr_scheduler_type="linear"), 5)
    for _ in range(9):
        opt.step()
        sched.step()
    assert opt.param_groups[0]["lr"] == 0.0                        # never negative past the horizon


def test_cosine_midpoint_floor_and_t_max():
    lrs = _run("cosine", steps=41, lr=1.0, eta_min=0.2)
    assert lrs[0] == pytest.approx(1.0) and lrs[20] == pytest.approx(0.2 + 0.8 * 0.5 * (1 + math.cos(math.pi * 20 / 41)))
    assert min(lrs) >= 0.2
    short = _run("cosine", steps=30, lr=1.0, T_max=10)
    assert short[10] == pytest.approx(0.0, abs=1e-9) and short[25] == pytest.approx(0.0, abs=1e-9)   # clamped after T_max


def test_polynomial_power():
    lrs = _run("polynomial", steps=10, lr=1.0, power=2.0)
    assert lrs[5] == pytest.approx((1 - 5 / 10) ** 2) and lrs[0] == 1.0


def test_step_decay_drops_exactly_at_multiples():
    lrs = _run("step", steps=12, lr=1.0, step_size=4, gamma=0.5)
    assert lrs[:4] == [1.0] * 4 and lrs[4:8] == [0.5] * 4 and lrs[8:12] == [0.25] * 4
    warm = _run("step", steps=12, lr=1.0, step_size=4, gamma=0.5, warmup_steps=2)
    assert warm[:2] == pytest.approx([0.5, 1.0]) and warm[2:6] == [1.0] * 4 and warm[6] == 0.5     # decay counts from the end of warm-up


def test_onecycle_peaks_at_pct_start():
    lrs = _run("onecycle", steps=50, lr=0.1, max_lr=1.0, pct_start=0.2)
    peak = max(range(50), key=lambda i: lrs[i])
    assert abs(peak - 0.2 * 50) <= 1 and max(lrs) == pytest.approx(1.0, rel=1e-3) and lrs[-1] < lrs[0]


def test_unknown_scheduler_and_custom_registration():
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=1.0)
    cfg = K.LrSchedulerArguments()
    cfg.lr_scheduler_type = "not-registered"                       # bypass the dataclass check: the factory checks too
    with pytest.raises(ValueError, match="registered"):
        create_lr_scheduler(opt, cfg, 10)

    def halving(optimizer, config, total):
        return torch.optim.lr_scheduler.LambdaLR(optimizer, lambda s: 0.5 ** s)

    assert register_scheduler("unit_halving", halving) is halving
    cfg = K.LrSchedulerArguments(lr_scheduler_type="unit_halving")  # registered names pass the dataclass validator
    sched = create_lr_scheduler(opt, cfg, 10)
    opt.step()
    sched.step()
    assert opt.param_groups[0]["lr"] == 0.5
    from scaletorch_b200.trainer import lr_scheduler as L

    L._REGISTRY.pop("unit_halving")


def test_total_steps_falls_back_to_config_then_default():
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=1.0)
    cfg = K.ScaleTorchArguments(lr_scheduler_type="linear", total_train_steps=4)
    sched = create_lr_scheduler(opt, cfg)                           # horizon from config.total_train_steps
    for _ in range(4):
        opt.step()
        sched.step()
    assert opt.param_groups[0]["lr"] == 0.0


def test_help_renders():
    """argparse %-expands help strings: a stray '%' in one of them breaks ``--help`` for every flag."""
    from transformers import HfArgumentParser

    text = HfArgumentParser(K.ScaleTorchArguments).format_help()
    assert "--tensor_parallel_size" in text and "--cp_zigzag" in text


def test_local_config_fast_path_agrees_with_autoconfig(tmp_path, monkeypatch):
    """Vendored configs are read directly (no ``AutoConfig`` import chain); every field the models read must equal what
    ``AutoConfig`` would have produced, and an incomplete ``config.json`` must still go through ``AutoConfig``."""
    import json

    from scaletorch_b200.trainer import config as K

    names = sorted(os.listdir(K._CONFIG_DIR))
    assert {"Qwen3-0.6B", "Qwen3-8B", "Qwen3-30B-A3B", "tiny-llama"} <= set(names)
    fields = ("model_type", "hidden_size", "intermediate_size", "num_hidden_layers", "num_attention_heads", "num_key_value_heads",
              "vocab_size", "max_position_embeddings", "rms_norm_eps", "tie_word_embeddings", "head_dim", "num_experts",
              "num_experts_per_tok", "moe_intermediate_size", "norm_topk_prob", "decoder_sparse_step", "mlp_only_layers",
              "router_aux_loss_coef", "eos_token_id")
    for name in names:
        fast = K.resolve_model_config(name)
        assert isinstance(fast, K.LocalModelConfig), name
        monkeypatch.setenv("SCALETORCH_B200_HF_AUTOCONFIG", "1")
        slow = K.resolve_model_config(name)
        monkeypatch.delenv("SCALETORCH_B200_HF_AUTOCONFIG")
        assert not isinstance(slow, K.LocalModelConfig)
        for f in fields:
            a, b = getattr(fast, f, None), getattr(slow, f, None)
            if f == "head_dim" and a is None:
                a = fast.hidden_size // fast.num_attention_heads
            assert a == b, (name, f, a, b)
        theta = getattr(slow, "rope_theta", None) or (getattr(slow, "rope_parameters", None) or {}).get("rope_theta")
        assert float(fast.rope_theta) == float(theta), name
        assert fast.to_dict()["hidden_size"] == fast.hidden_size and fast.get("no_such_key", 7) == 7
    # a file that leaves a family field to the class default is NOT taken by the fast path
    d = json.load(open(os.path.join(K._CONFIG_DIR, "tiny-qwen3", "config.json")))
    d.pop("rms_norm_eps")
    (tmp_path / "config.json").write_text(json.dumps(d))
    cfg = K.resolve_model_config(str(tmp_path))
    assert not isinstance(cfg, K.LocalModelConfig) and cfg.rms_norm_eps == 1e-6
    # unknown families too
    d["model_type"] = "gpt2"
    (tmp_path / "config.json").write_text(json.dumps(d))
    assert K._complete_local_config(str(tmp_path)) is None
