"""Config dataclasses, LR schedulers, metrics line, optimizer / flat store, data loader, monitor (CPU)."""
import re

import pytest
import torch

from scaletorch_b200.trainer.config import (LrSchedulerArguments, OptimizerArguments, ParallelArguments,
                                            ScaleTorchArguments, TrainingArguments)


def test_config_defaults_match_reference():
    c = ScaleTorchArguments()
    assert c.model_name_or_path == "facebook/opt-125m" and c.dataset_name == "wikitext2"
    assert c.sequence_length == 1024 and c.seed == 1 and c.learning_rate == 1e-3
    assert c.lr_scheduler_type == "linear" and c.max_grad_norm == 1.0 and c.save_frequency == 300
    assert c.use_fused_adam and c.pipeline_parallel_engine == "1f1b" and c.backend == "nccl"
    assert c.micro_batch_size == c.batch_size == 64
    assert c.global_batch_size == 64 and c.global_batch_size_token == 64 * 1024


def test_config_validation():
    with pytest.raises(ValueError):
        ParallelArguments(tensor_parallel_size=0)
    with pytest.raises(ValueError):
        ParallelArguments(expert_parallel_size=0)
    with pytest.raises(ValueError):
        ParallelArguments(pipeline_parallel_engine="zero-bubble")
    with pytest.raises(ValueError):
        ParallelArguments(backend="hccl")
    with pytest.raises(ValueError):
        LrSchedulerArguments(lr_scheduler_type="nope")
    with pytest.raises(ValueError):
        LrSchedulerArguments(lr_scheduler_type="step", gamma=1.5)
    with pytest.raises(ValueError):
        OptimizerArguments(optimizer_type="adagrad")
    with pytest.raises(ValueError):
        TrainingArguments(gradient_accumulation_steps=0)
    with pytest.raises(ValueError):
        ScaleTorchArguments(sequence_length=1000, context_parallel_size=3)
    c = ScaleTorchArguments(data_parallel_size=2, expert_parallel_size=2, micro_batch_size=3, gradient_accumulation_steps=2,
                            tensor_parallel_size=2)
    assert c.global_batch_size == 2 * 2 * 3 * 2
    c.validate_world_size(8)
    with pytest.raises(ValueError):
        c.validate_world_size(4)


def test_hf_argument_parser_roundtrip():
    from transformers import HfArgumentParser

    (c,) = HfArgumentParser(ScaleTorchArguments).parse_args_into_dataclasses(
        args=["--model_name_or_path", "tiny-qwen3", "--tensor_parallel_size", "2", "--sequence_parallel", "True",
              "--betas", "0.9", "0.95", "--micro_batch_size", "4", "--synthetic_data", "True"])
    assert c.tensor_parallel_size == 2 and c.sequence_parallel and tuple(c.betas) == (0.9, 0.95)
    assert c.num_hidden_layers == 2  # filled from the vendored config


@pytest.mark.parametrize("kind", ["linear", "cosine", "polynomial", "step", "onecycle", "constant"])
def test_lr_schedulers(kind):
    from scaletorch_b200.trainer.lr_scheduler import create_lr_scheduler, register_scheduler

    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([p], lr=1.0)
    cfg = LrSchedulerArguments(lr_scheduler_type=kind, warmup_steps=5, eta_min=0.1, step_size=10, gamma=0.5, power=2.0)
    cfg.total_train_steps = 50
    sched = create_lr_scheduler(opt, cfg, 50)
    lrs = []
    for _ in range(50):
        lrs.append(opt.param_groups[0]["lr"])
        opt.step()
        sched.step()
    if kind != "onecycle":
        assert lrs[0] == pytest.approx(0.2) and lrs[4] == pytest.approx(1.0)  # linear warm-up
        assert all(b <= a + 1e-9 for a, b in zip(lrs[5:], lrs[6:]))           # non-increasing afterwards
    if kind == "cosine":
        assert min(lrs) >= 0.1 - 1e-6
    if kind == "constant":
        assert lrs[-1] == 1.0

    @register_scheduler("my_flat")
    def _flat(optimizer, config, total):
        return torch.optim.lr_scheduler.LambdaLR(optimizer, lambda s: 0.5)

    cfg2 = LrSchedulerArguments(lr_scheduler_type="my_flat")
    assert create_lr_scheduler(opt, cfg2, 10) is not None


def test_metrics_line_is_regex_parsable_like_the_reference_scripts():
    from scaletorch_b200.trainer.metrics import format_metrics_line

    line = format_metrics_line(12, 3.14159, 1e-3, 0.5, 8192, 123456.0, 15432.0, 98304, None, 41.5, 21.3)
    assert re.search(r"Tokens/s/GPU:\s*([\d.]+[KMBT]?)", line).group(1) == "15.43K"
    assert re.search(r"MFU:\s*([\d.]+)%", line).group(1) == "41.50"
    assert re.search(r"Loss:\s*([\d.]+)", line).group(1) == "3.1416"
    assert line.count(" | ") == 9


def test_flat_store_and_fused_adamw_match_torch_adamw():
    from scaletorch_b200.runtime.flat_params import FlatParamStore
    from scaletorch_b200.trainer.optimizer import FusedAdamW

    torch.manual_seed(0)
    m1 = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Linear(16, 4))
    m2 = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Linear(16, 4))
    m2.load_state_dict(m1.state_dict())
    store = FlatParamStore(m1)
    assert store.total_numel >= sum(p.numel() for p in m1.parameters())
    assert all(p.data_ptr() >= store.param_flat.data_ptr() for p in m1.parameters())
    o1 = FusedAdamW(store, lr=1e-2, betas=(0.9, 0.95), weight_decay=0.1, max_grad_norm=0.5)
    o2 = torch.optim.AdamW(m2.parameters(), lr=1e-2, betas=(0.9, 0.95), weight_decay=0.1)
    for _ in range(3):
        x = torch.randn(5, 8)
        o1.zero_grad()
        m1(x).pow(2).sum().backward()          # grads arrive through the fallback hook into main_grad
        o1.step()
        o2.zero_grad()
        m2(x).pow(2).sum().backward()
        torch.nn.utils.clip_grad_norm_(m2.parameters(), 0.5)
        o2.step()
    for a, b in zip(m1.parameters(), m2.parameters()):
        assert torch.allclose(a, b, atol=1e-5), (a - b).abs().max()
    assert o1.grad_norm() > 0
    sd = o1.state_dict()
    o1.load_state_dict(sd)


def test_microbatch_dataloader_cp_slicing_and_shift():
    from scaletorch_b200.data import MicroBatchDataLoader

    dl = MicroBatchDataLoader(2, 16, synthetic=True, vocab_size=100, num_samples=8, gradient_accumulation_steps=2)
    b = next(dl)
    assert b["input_ids"].shape == (2, 16) and b["target_ids"].shape == (2, 16)
    assert torch.equal(b["input_ids"][:, 1:], b["target_ids"][:, :-1])
    assert torch.equal(b["position_ids"][0], torch.arange(16))
    assert dl.global_batch_size == 4
    for _ in range(10):   # wraps around the 8-sample dataset
        next(dl)
    # manual CP slice
    dl.cp_world_size, dl.cp_rank, dl.sequence_length_per_rank = 2, 1, 8
    c = dl.collate_batch([{"input_ids": torch.arange(17)}])
    assert torch.equal(c["input_ids"][0], torch.arange(8, 16)) and torch.equal(c["target_ids"][0], torch.arange(9, 17))
    assert torch.equal(c["position_ids"][0], torch.arange(8, 16))


def test_concat_chunk_strategy_with_fake_tokenizer():
    from scaletorch_b200.data.dataset import get_tokenize_strategy, register_tokenize_strategy

    class Tok:
        def __call__(self, texts, **kw):
            return {"input_ids": [[ord(c) % 50 for c in t] for t in texts]}

    out = get_tokenize_strategy("concat_chunk")({"text": ["abcdefgh", "ijklmnopq"]}, Tok(), 4)
    assert all(len(c) == 5 for c in out["input_ids"]) and len(out["input_ids"]) == 3
    with pytest.raises(KeyError):
        get_tokenize_strategy("nope")
    register_tokenize_strategy("mine", lambda ex, tok, S, col="text": {"input_ids": []})
    assert get_tokenize_strategy("mine") is not None


def test_utils_misc_and_monitor(tmp_path):
    from types import SimpleNamespace

    from scaletorch_b200.utils import PerformanceMonitor, get_mfu, to_readable_format
    from scaletorch_b200.utils.device import get_theoretical_flops, register_device_flops
    from scaletorch_b200.utils.misc import flops_per_token

    assert to_readable_format(1500) == "1.50K" and to_readable_format(2.5e9) == "2.50B" and to_readable_format(12) == "12.00"
    cfg = SimpleNamespace(num_hidden_layers=28, num_attention_heads=16, head_dim=128, hidden_size=1024, max_position_embeddings=4096)
    f = flops_per_token(596e6, 28, 16, 128, 2048)
    assert f == 6 * 596e6 + 12 * 28 * 16 * 128 * 2048
    assert get_mfu(1000.0, 596e6, cfg, theoretical_flops=1e15, seq_len=2048) == pytest.approx(1000 * f / 1e15 * 100)
    register_device_flops("NoSuchDevice", 1e12)
    assert get_theoretical_flops() > 0
    mon = PerformanceMonitor(config=SimpleNamespace(a=1), log_dir=str(tmp_path))
    mon.start()
    for _ in range(3):
        mon.start_iteration(100)
        s = mon.end_iteration()
        assert s["tokens"] == 100
    assert mon.get_average_stats()["num_iterations"] == 3
    assert mon.save_stats("perf.json").endswith("perf.json")
    with pytest.raises(RuntimeError):
        mon.end_iteration()
