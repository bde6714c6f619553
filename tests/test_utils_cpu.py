"""``scaletorch_b200.utils`` on CPU: seeding, number formatting, MFU arithmetic, parameter counting, meta-tensor guard,
loss averaging, device helpers + the peak-FLOPS registry, the performance monitor, path / logger helpers.
Reference suites: ``tests/utils/test_utils.py``, ``test_device.py``, ``test_device_monitor.py``.
"""
import json
import logging
import os
import random
import time
from types import SimpleNamespace

import numpy as np
import pytest
import torch
import torch.nn as nn

from scaletorch_b200.utils import device as dev
from scaletorch_b200.utils import misc
from scaletorch_b200.utils.monitor import PerformanceMonitor


# --------------------------------------------------------------------------------------------- seeding
def test_set_all_seed_reproducible_and_distinct():
    def draw():
        return random.random(), float(np.random.rand()), float(torch.rand(1))

    misc.set_all_seed(42)
    a = draw()
    misc.set_all_seed(42)
    assert draw() == a
    misc.set_all_seed(43)
    assert draw() != a
    misc.set_all_seed(0)                                         # zero is a valid seed
    misc.set_all_seed(42, rank_offset=7)                         # model-parallel shards: same seed, different stream
    assert draw() != a
    misc.set_all_seed(2**40)                                     # numpy's 32-bit limit is handled


@pytest.mark.parametrize("bad,err", [(-1, ValueError), (1.5, TypeError), ("7", TypeError), (None, TypeError), (True, TypeError)])
def test_set_all_seed_rejects(bad, err):
    with pytest.raises(err):
        misc.set_all_seed(bad)


def test_set_all_seed_deterministic_flag():
    misc.set_all_seed(1, deterministic=True)
    assert torch.backends.cudnn.deterministic and not torch.backends.cudnn.benchmark
    torch.backends.cudnn.deterministic = False


# --------------------------------------------------------------------------------------------- formatting
@pytest.mark.parametrize("num,want", [(0, "0.00"), (999, "999.00"), (1500, "1.50K"), (2_500_000, "2.50M"),
                                      (7.5e9, "7.50B"), (1.25e12, "1.25T"), (-3200, "-3.20K"), (1234.5678, "1.23K")])
def test_to_readable_format(num, want):
    assert misc.to_readable_format(num) == want


def test_to_readable_format_precision_and_errors():
    assert misc.to_readable_format(1_234_567, precision=0) == "1M"
    assert misc.to_readable_format(1_234_567, precision=4) == "1.2346M"
    assert misc.to_readable_format(np.int64(2048)) == "2.05K"
    with pytest.raises(ValueError):
        misc.to_readable_format(1, precision=-1)
    with pytest.raises(TypeError):
        misc.to_readable_format("12")
    with pytest.raises(TypeError):
        misc.to_readable_format(None)


# --------------------------------------------------------------------------------------------- MFU
_CFG = SimpleNamespace(num_hidden_layers=4, num_attention_heads=8, hidden_size=512, head_dim=64, max_position_embeddings=1024)


def test_flops_per_token_formula():
    assert misc.flops_per_token(1e6, 4, 8, 64, 1024) == 6e6 + 12 * 4 * 8 * 64 * 1024


def test_get_mfu_basic_and_explicit_peak():
    f = misc.flops_per_token(1e9, 4, 8, 64, 1024)
    assert misc.get_mfu(1000.0, 1e9, _CFG, theoretical_flops=1e15) == pytest.approx(1000.0 * f / 1e15 * 100)
    # seq_len argument overrides max_position_embeddings
    f2 = misc.flops_per_token(1e9, 4, 8, 64, 256)
    assert misc.get_mfu(1000.0, 1e9, _CFG, theoretical_flops=1e15, seq_len=256) == pytest.approx(1000.0 * f2 / 1e15 * 100)
    # head_dim falls back to hidden_size / heads
    cfg = SimpleNamespace(num_hidden_layers=4, num_attention_heads=8, hidden_size=512, max_position_embeddings=1024)
    assert misc.get_mfu(1000.0, 1e9, cfg, theoretical_flops=1e15) == pytest.approx(1000.0 * f / 1e15 * 100)


@pytest.mark.parametrize("tps,n,peak", [(0, 1e9, 1e15), (-5, 1e9, 1e15), (100, 0, 1e15), (100, -1, 1e15), (100, 1e9, 0),
                                        (100, 1e9, -1), (None, 1e9, 1e15)])
def test_get_mfu_degenerate_inputs_give_zero(tps, n, peak):
    assert misc.get_mfu(tps, n, _CFG, theoretical_flops=peak) == 0.0


def test_get_mfu_uses_env_peak(monkeypatch):
    monkeypatch.setenv("SCALETORCH_DEVICE_FLOPS", "2e15")
    a = misc.get_mfu(1000.0, 1e9, _CFG)
    monkeypatch.setenv("SCALETORCH_DEVICE_FLOPS", "1e15")
    assert misc.get_mfu(1000.0, 1e9, _CFG) == pytest.approx(2 * a)


def test_active_params_dense_is_identity_and_moe_counts_topk():
    assert misc.active_params(123.0, _CFG) == 123.0
    moe = SimpleNamespace(num_hidden_layers=2, hidden_size=64, num_experts=8, num_experts_per_tok=2, moe_intermediate_size=32,
                          decoder_sparse_step=1, mlp_only_layers=[])
    expert_total = 2 * 8 * 3 * 64 * 32
    assert misc.active_params(1e6, moe) == pytest.approx(1e6 - expert_total * (1 - 2 / 8))
    moe.mlp_only_layers = [0]                                    # only one sparse layer left
    assert misc.active_params(1e6, moe) == pytest.approx(1e6 - (expert_total / 2) * 0.75)


# --------------------------------------------------------------------------------------------- parameter counting
class _Toy(nn.Module):
    def __init__(self):
        super().__init__()
        self.embedding = nn.Embedding(10, 4)
        self.q_proj = nn.Linear(4, 4, bias=False)
        self.norm = nn.LayerNorm(4)
        self.final_proj = nn.Linear(4, 10, bias=False)
        self.final_proj.weight = self.embedding.weight            # tied: counted once


def test_get_num_params_without_process_group():
    m = _Toy()
    assert misc.get_num_params(m) == 40 + 16 + 8
    assert misc.get_num_params(nn.Sequential()) == 0


def test_is_tp_sharded_name():
    assert misc.is_tp_sharded_name("layers.0.attention.qkv_proj.weight")
    assert misc.is_tp_sharded_name("embedding.weight") and misc.is_tp_sharded_name("layers.3.mlp.down_proj.weight")
    assert not misc.is_tp_sharded_name("layers.0.attention.q_norm.weight")
    assert not misc.is_tp_sharded_name("layers.0.mlp.router.gate.weight")
    assert not misc.is_tp_sharded_name("layers.0.mlp.experts.gate_up_proj")


def test_assert_no_meta_tensors():
    misc.assert_no_meta_tensors(nn.Linear(2, 2))
    with torch.device("meta"):
        m = nn.Linear(2, 2)
    with pytest.raises(RuntimeError, match="meta"):
        misc.assert_no_meta_tensors(m)
    ok = nn.Linear(2, 2)
    ok.register_buffer("b", torch.empty(1, device="meta"))
    with pytest.raises(RuntimeError, match="b"):
        misc.assert_no_meta_tensors(ok)


def test_average_loss_single_process():
    cpu = torch.device("cpu")
    assert misc.average_loss_across_dp_cp_ranks(2.5, cpu) == 2.5
    assert misc.average_loss_across_dp_cp_ranks(None, cpu) == 0.0
    assert misc.average_loss_across_dp_cp_ranks(torch.tensor(1.25, dtype=torch.bfloat16), cpu) == 1.25
    t = misc.average_loss_across_dp_cp_ranks(torch.tensor([3.0]), cpu, as_tensor=True)
    assert isinstance(t, torch.Tensor) and t.shape == (1,) and t.dtype == torch.float32
    with pytest.raises(TypeError):
        misc.average_loss_across_dp_cp_ranks("nan", cpu)


def test_rank_print_lock_and_mute(capsys, tmp_path, monkeypatch):
    monkeypatch.setenv("SCALETORCH_PRINT_LOCK", str(tmp_path / "lock"))
    misc.rank_print("hello", 3)
    misc.rank_print("muted", is_print_rank=False)
    out = capsys.readouterr().out
    assert "hello 3" in out and "muted" not in out
    monkeypatch.setenv("SCALETORCH_PRINT_LOCK", str(tmp_path / "no" / "such" / "dir" / "lock"))
    misc.rank_print("still printed")                             # lock file cannot be created: print anyway
    assert "still printed" in capsys.readouterr().out


# --------------------------------------------------------------------------------------------- device helpers
def test_device_helpers_on_cpu(monkeypatch):
    if torch.cuda.is_available():
        pytest.skip("CPU expectations")
    assert dev.get_device_type() == "cpu" and not dev.is_accelerator_available() and dev.get_dist_backend() == "gloo"
    assert dev.get_visible_devices_keyword() == "CUDA_VISIBLE_DEVICES"
    assert dev.get_current_device() == torch.device("cpu") and dev.get_current_device(use_cpu=True) == torch.device("cpu")
    assert dev.get_device_count() == 0
    assert dev.memory_allocated() == dev.memory_reserved() == dev.max_memory_allocated() == 0
    dev.synchronize(), dev.empty_cache(), dev.reset_peak_memory_stats(), dev.set_device(0), dev.manual_seed_all(3)
    assert dev.is_bf16_supported()


def test_dist_info_from_env(monkeypatch):
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    assert dev.get_dist_info() == (0, 1, 0)
    monkeypatch.setenv("RANK", "5"), monkeypatch.setenv("WORLD_SIZE", "8"), monkeypatch.setenv("LOCAL_RANK", "1")
    assert dev.get_dist_info() == (5, 8, 1)


def test_theoretical_flops_env_override_and_registry(monkeypatch):
    monkeypatch.setenv("SCALETORCH_DEVICE_FLOPS", "1.5e15")
    assert dev.get_theoretical_flops() == 1.5e15
    monkeypatch.setenv("SCALETORCH_DEVICE_FLOPS", "garbage")      # unusable override: fall through, never raise
    assert dev.get_theoretical_flops() > 0
    monkeypatch.setenv("SCALETORCH_DEVICE_FLOPS", "-3")
    assert dev.get_theoretical_flops() > 0
    monkeypatch.delenv("SCALETORCH_DEVICE_FLOPS")
    assert dev.get_theoretical_flops() > 0
    with pytest.raises(ValueError):
        dev.register_device_flops("X", 0)
    n = len(dev._DEVICE_FLOPS)
    dev.register_device_flops("UnitTestGPU", 1e12)
    assert dev._DEVICE_FLOPS[0] == ("UnitTestGPU", 1e12) and len(dev._DEVICE_FLOPS) == n + 1
    dev._DEVICE_FLOPS.pop(0)
    names = [k for k, _ in dev._DEVICE_FLOPS]
    assert "B200" in names and dict(dev._DEVICE_FLOPS)["B200"] == 2.25e15


def test_measured_peaks_is_a_dict():
    assert isinstance(dev.get_measured_peaks(), dict)


# --------------------------------------------------------------------------------------------- performance monitor
def test_monitor_lifecycle_and_averages(tmp_path):
    mon = PerformanceMonitor(config=SimpleNamespace(work_dir=str(tmp_path), lr=1e-3, obj=object()))
    with pytest.raises(RuntimeError):
        mon.end_iteration()
    with pytest.raises(RuntimeError):
        mon.end_iteration_deferred()
    assert mon.get_average_stats() == {} and mon.save_stats() is None
    mon.print_summary()                                          # no data: must not raise
    mon.start()
    for i in range(3):
        mon.start_iteration(tokens_processed=100 * (i + 1))
        time.sleep(0.002)
        s = mon.end_iteration()
        assert s["iteration"] == i + 1 and s["tokens"] == 100 * (i + 1) and s["iteration_time"] > 0
        assert s["tokens_per_second"] == pytest.approx(s["tokens"] / s["iteration_time"])
    mon.start_iteration()
    h = mon.end_iteration_deferred()                             # resolved later (lagged logging)
    mon.start_iteration(50)
    mon.end_iteration(tokens_processed=70)                       # token count may be given at the end
    s = mon.resolve_iteration(h)
    assert s["iteration"] == 5 and mon.total_tokens == 100 + 200 + 300 + 70
    avg = mon.get_average_stats()
    assert avg["num_iterations"] == 5 and avg["total_tokens"] == 670 and avg["total_time"] > 0
    assert avg["avg_tokens"] == pytest.approx(670 / 5)
    assert mon.get_average_stats(skip=3)["num_iterations"] == 2  # skip warm-up iterations
    mon.log_iteration_stats(s)
    mon.print_summary()
    path = mon.save_stats("perf.json")
    assert path == str(tmp_path / "perf.json")
    blob = json.load(open(path))
    assert len(blob["iterations"]) == 5 and blob["config"]["lr"] == 1e-3 and isinstance(blob["config"]["obj"], str)
    auto = mon.save_stats()
    assert os.path.basename(auto).startswith("performance_logs_0_") and os.path.isfile(auto)
    mon.close()
    mon.close()                                                  # idempotent


def test_monitor_history_is_bounded(monkeypatch):
    import scaletorch_b200.utils.monitor as M

    monkeypatch.setattr(M, "MAX_HISTORY", 10)
    monkeypatch.setattr(M, "TRIM_TO", 4)
    mon = PerformanceMonitor()
    for _ in range(11):
        mon._record(0.01, 1)
    assert len(mon.history) == 4 and mon.history[-1]["iteration"] == 11 and mon.total_tokens == 11


# --------------------------------------------------------------------------------------------- path / logger / env
def test_path_helpers(tmp_path):
    from scaletorch_b200.utils.path import check_file_exist, mkdir_or_exist

    f = tmp_path / "a.txt"
    f.write_text("x")
    check_file_exist(str(f))
    with pytest.raises(FileNotFoundError, match="missing.bin"):
        check_file_exist(str(tmp_path / "missing.bin"))
    with pytest.raises(FileNotFoundError, match="custom"):
        check_file_exist(str(tmp_path / "missing.bin"), msg_tmpl="custom {}")
    d = tmp_path / "x" / "y"
    mkdir_or_exist(str(d))
    mkdir_or_exist(str(d))
    assert d.is_dir()
    mkdir_or_exist("")                                           # empty name is ignored


def test_logger_and_outdir(tmp_path):
    from scaletorch_b200.utils.logger_utils import get_logger, get_outdir

    log = tmp_path / "run.log"
    lg = get_logger("unit.test.logger", log_file=str(log), log_level=logging.DEBUG)
    assert get_logger("unit.test.logger") is lg                  # configured once, no duplicate handlers
    n = len(lg.handlers)
    get_logger("unit.test.logger", log_file=str(log))
    assert len(lg.handlers) == n
    lg.info("hello %d", 7)
    for h in lg.handlers:
        h.flush()
    assert "hello 7" in log.read_text()
    out = get_outdir(str(tmp_path), "exp")
    assert os.path.isdir(out) and out.endswith("exp")
    out2 = get_outdir(str(tmp_path), "exp", inc=True)            # existing directory: numbered sibling
    assert out2 != out and os.path.isdir(out2)


def test_system_info_keys():
    from scaletorch_b200.utils import get_system_info

    info = get_system_info()
    assert isinstance(info, dict) and any("torch" in k.lower() for k in info) and any("python" in k.lower() for k in info)
