"""Small-unit tests mirroring the reference's mock-based suite: communicator state machine, pipeline layer
distribution, HF name mapping, device helpers, benchmark-script plumbing."""
import importlib.util
import os
import sys
from types import SimpleNamespace

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, path))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


# ------------------------------------------------------------------ context-parallel communicator
def test_context_communicator_state_machine_single_rank():
    from scaletorch_b200.parallel.context_parallel.cp_comms import ContextCommunicator

    c = ContextCommunicator("t")
    x = torch.arange(6.0).view(2, 3)
    with pytest.raises(RuntimeError, match="wait called before commit"):
        c.wait()
    r = c.send_recv(x)
    assert torch.equal(r, x)                      # world size 1: the "ring" delivers the tensor to ourselves
    c.commit()
    with pytest.raises(RuntimeError, match="commit called twice"):
        c.commit()
    with pytest.raises(RuntimeError, match="after commit"):
        c.send_recv(x)
    c.wait()
    c.send_recv(x)                                 # reusable after wait()
    c.commit()
    c.wait()


# ------------------------------------------------------------------ pipeline layer split
@pytest.mark.parametrize("L,pp", [(28, 2), (36, 4), (10, 4), (7, 8), (48, 3)])
def test_distribute_layers_is_a_contiguous_partition(L, pp):
    from scaletorch_b200.parallel.pipeline_parallel.pipeline_parallel import distribute_layers

    parts = [distribute_layers(L, pp, r) for r in range(pp)]
    assert [i for p in parts for i in p] == list(range(L))
    sizes = [len(p) for p in parts]
    assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)     # the first L % pp stages get one extra


# ------------------------------------------------------------------ HF <-> local names
def test_hf_name_mapping_covers_every_parameter():
    from scaletorch_b200.models import Qwen3, Qwen3MoE
    from scaletorch_b200.parallel.context import ParallelContext
    from scaletorch_b200.utils.checkpoint import InitializationManager

    base = dict(vocab_size=256, hidden_size=64, num_attention_heads=2, num_key_value_heads=1, head_dim=32,
                max_position_embeddings=64, num_hidden_layers=2, intermediate_size=128, rms_norm_eps=1e-6, rope_theta=1e4,
                tie_word_embeddings=False)
    dense = SimpleNamespace(model_type="qwen3", **base)
    m = Qwen3(dense, ParallelContext())
    im = InitializationManager(m, dense)
    names = [h for n, _ in m.named_parameters() for h in im.hf_names_for_local(n)]
    assert "model.layers.1.self_attn.q_norm.weight" in names and "lm_head.weight" in names
    assert [n for n in names if "layers.0.self_attn" in n and "proj" in n] == [
        "model.layers.0.self_attn.q_proj.weight", "model.layers.0.self_attn.k_proj.weight",
        "model.layers.0.self_attn.v_proj.weight", "model.layers.0.self_attn.o_proj.weight"]
    assert len(names) == len(set(names))
    moe_cfg = SimpleNamespace(model_type="qwen3_moe", num_experts=4, num_experts_per_tok=2, moe_intermediate_size=32,
                              norm_topk_prob=True, decoder_sparse_step=1, mlp_only_layers=[], router_aux_loss_coef=0.0, **base)
    mm = Qwen3MoE(moe_cfg, ParallelContext())
    imm = InitializationManager(mm, moe_cfg)
    got = imm.hf_names_for_local("decoder_layers.1.moe.experts.gate_up_proj")
    assert got[:2] == ["model.layers.1.mlp.experts.0.gate_proj.weight", "model.layers.1.mlp.experts.0.up_proj.weight"]
    assert len(got) == 8 and imm.hf_names_for_local("decoder_layers.0.moe.router.gate.weight") == ["model.layers.0.mlp.gate.weight"]
    with pytest.raises(KeyError):
        im.hf_names_for_local("decoder_layers.0.nonexistent.weight")


# ------------------------------------------------------------------ device helpers
def test_device_helpers_and_flops_registry(monkeypatch):
    from scaletorch_b200.utils import device as D

    assert D.get_device_type() in ("cuda", "cpu")
    assert D.get_dist_backend() in ("nccl", "gloo")
    assert D.get_current_device(use_cpu=True).type == "cpu"
    monkeypatch.setenv("SCALETORCH_DEVICE_FLOPS", "3.5e14")
    assert D.get_theoretical_flops() == 3.5e14
    monkeypatch.setenv("SCALETORCH_DEVICE_FLOPS", "garbage")
    assert D.get_theoretical_flops() > 0
    with pytest.raises(ValueError):
        D.register_device_flops("X", 0.0)
    n = len(D._DEVICE_FLOPS)
    D.register_device_flops("MyAccel", 1e15)
    assert D._DEVICE_FLOPS[0] == ("MyAccel", 1e15) and len(D._DEVICE_FLOPS) == n + 1
    D._DEVICE_FLOPS.pop(0)
    assert any(sub == "B200" and f == 2.25e15 for sub, f in D._DEVICE_FLOPS)
    assert isinstance(D.get_measured_peaks(), dict)


def test_active_params_for_moe_mfu():
    import json

    from scaletorch_b200.utils.misc import active_params

    cfg = SimpleNamespace(**json.load(open(os.path.join(ROOT, "configs", "Qwen3-30B-A3B", "config.json"))))
    act = active_params(30.5e9, cfg)
    assert 3.0e9 < act < 3.6e9                         # "A3B"
    dense = SimpleNamespace(num_hidden_layers=2, hidden_size=64)
    assert active_params(123.0, dense) == 123.0


# ------------------------------------------------------------------ scripts
def test_benchmark_comprehensive_matrix_and_commands():
    bc = _load("scripts/benchmark_comprehensive.py", "bench_comp")
    cfgs = bc.matrix()
    assert len(cfgs) >= 50 and len({c.label for c in cfgs}) == len(cfgs)
    assert all(8 % c.world_min() == 0 for c in cfgs)
    c = next(c for c in cfgs if c.label == "Qwen3-8B SP-TP2-CP2 S8192")
    cmd = bc.command(c, 8, "ours", 10, 3, 29999)
    assert "--nproc-per-node=8" in cmd and cmd[cmd.index("--tp") + 1] == "2" and cmd[cmd.index("--cp") + 1] == "2"
    assert "--sequence_parallel" in cmd and "--gradient_checkpointing" not in cmd
    one = bc.command(bc.Cfg("x", "Qwen3-0.6B"), 1, "reference", 5, 3, 1)
    assert "torch.distributed.run" not in one and one[one.index("--impl") + 1] == "reference"
    md = bc.markdown([{"label": "a", "impl": "ours", "gpus": 8, "status": "ok",
                       "result": {"value": 800000.0, "ms_per_step": 40.0, "mfu_pct_of_nominal_2.25PF": 20.0}},
                      {"label": "b", "impl": "reference", "gpus": 8, "status": "failed"}])
    assert "| a | ours | 8 | 800,000 | 100,000 | 40.0 | 20.0 | ok |" in md and "| b | reference | 8 | – |" in md


def test_group_nodes_by_rack():
    gn = _load("scripts/group_nodes.py", "group_nodes")
    racks = gn.group(["10.0.0.1\tr1", "# comment", "", "10.0.0.2 r2", "10.0.0.3\tr1", "10.0.0.4"])
    assert racks[0] == ("r1", ["10.0.0.1", "10.0.0.3"]) and dict(racks)["unknown"] == ["10.0.0.4"]


def test_bench_config_dict_names_the_layout():
    bench = _load("bench.py", "bench_mod")
    args = SimpleNamespace(tp=2, cp=2, pp=1, ep=1, micro_batch_size=1, grad_accum=1, seq_len=8192, model="Qwen3-8B",
                           sequence_parallel=True, gradient_checkpointing=False)
    dp, cfg = bench.build_config_dict(args, 8)
    assert dp == 2 and cfg["parallelism"] == "tp2cp2dp2" and cfg["global_batch"] == 2 and "L2" in cfg["l2_policy"]
    args.ep, args.tp, args.cp = 8, 1, 1
    dp, cfg = bench.build_config_dict(args, 8)
    assert dp == 1 and cfg["parallelism"] == "ep8dp1" and cfg["global_batch"] == 8     # EP ranks read distinct data


def test_graphed_step_support_matrix(monkeypatch):
    """Whole-step CUDA-graph replay is offered for the layouts validated on hardware; TP / EP can be opted out."""
    import scaletorch_b200.trainer.graphed as G
    from scaletorch_b200.trainer.optimizer import FusedAdamW

    opt = FusedAdamW.__new__(FusedAdamW)
    cfg = SimpleNamespace(fused_comm=True)
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)

    def pgm(**kw):
        base = dict(pp_world_size=1, ep_world_size=1, tp_world_size=1, world_size=8)
        base.update(kw)
        m = SimpleNamespace(**base)
        return m

    monkeypatch.setattr(G, "pgm", pgm())
    assert G.GraphedTrainStep.supported(cfg, opt)                                   # DP / CP
    assert not G.GraphedTrainStep.supported(cfg, torch.optim.SGD([torch.nn.Parameter(torch.zeros(1))], lr=0.1))
    monkeypatch.delenv("SCALETORCH_B200_GRAPH_MP", raising=False)
    monkeypatch.setattr(G, "pgm", pgm(pp_world_size=2))
    assert not G.GraphedTrainStep.supported(cfg, opt)                               # pipeline schedules stay eager
    for mp in (dict(ep_world_size=2), dict(tp_world_size=2)):                       # TP / EP: captured by default ...
        monkeypatch.setattr(G, "pgm", pgm(**mp))
        assert G.GraphedTrainStep.supported(cfg, opt), mp
        monkeypatch.setenv("SCALETORCH_B200_GRAPH_MP", "none")                      # ... unless switched off
        assert not G.GraphedTrainStep.supported(cfg, opt), mp
        monkeypatch.setenv("SCALETORCH_B200_GRAPH_MP", "tp" if "ep_world_size" in mp else "ep")
        assert not G.GraphedTrainStep.supported(cfg, opt), mp                       # only the other layout allowed
        monkeypatch.delenv("SCALETORCH_B200_GRAPH_MP")
    monkeypatch.setattr(G, "pgm", pgm())
    assert not G.GraphedTrainStep.supported(SimpleNamespace(fused_comm=False), opt)  # c10d collectives are not captured
    from scaletorch_b200.models import Qwen3MoE
    from scaletorch_b200.parallel.context import ParallelContext

    moe_cfg = SimpleNamespace(model_type="qwen3_moe", num_experts=4, num_experts_per_tok=2, moe_intermediate_size=32,
                              norm_topk_prob=True, decoder_sparse_step=1, mlp_only_layers=[], router_aux_loss_coef=0.0,
                              vocab_size=64, hidden_size=32, num_attention_heads=2, num_key_value_heads=1, head_dim=16,
                              max_position_embeddings=32, num_hidden_layers=1, intermediate_size=64, rms_norm_eps=1e-6,
                              rope_theta=1e4, tie_word_embeddings=False)
    moe = Qwen3MoE(moe_cfg, ParallelContext())
    assert G.GraphedTrainStep.supported(cfg, opt, moe)                              # MoE steps capture (side-stream warm-up)
    monkeypatch.setenv("SCALETORCH_B200_GRAPH_MP", "tp")
    assert not G.GraphedTrainStep.supported(cfg, opt, moe)
    monkeypatch.delenv("SCALETORCH_B200_GRAPH_MP")
    monkeypatch.setattr(torch.cuda, "is_available", lambda: False)
    assert not G.GraphedTrainStep.supported(cfg, opt)


def test_fused_tp_policy_and_cpu_predicates(monkeypatch):
    """The measured policy: fused TP kernels from 4 ranks up, overridable; CPU tensors never take the fused path."""
    from scaletorch_b200.parallel import fused as Fz

    monkeypatch.delenv("SCALETORCH_B200_FUSED_TP", raising=False)
    assert [Fz._policy("gemm_rs", w) for w in (1, 2, 3, 4, 8)] == [False, False, False, True, True]
    monkeypatch.setenv("SCALETORCH_B200_FUSED_TP", "all")
    assert Fz._policy("ag_gemm", 2)
    monkeypatch.setenv("SCALETORCH_B200_FUSED_TP", "none")
    assert not Fz._policy("gemm_ar", 8)
    monkeypatch.setenv("SCALETORCH_B200_FUSED_TP", "auto")
    assert not Fz._policy("gemm_ar", 2) and Fz._policy("gemm_ar", 4)
    x = torch.zeros(256, 64, dtype=torch.bfloat16)
    for pred in (Fz.ag_gemm_available, Fz.gemm_rs_available, Fz.allreduce_available):
        assert not pred(x, None)                                  # not CUDA: the c10d path serves it
    assert not Fz.gemm_ar_available(x, 128, None)


def test_ep_default_capacity_is_dropless(monkeypatch):
    from scaletorch_b200.parallel.expert_parallel import fused_ep

    monkeypatch.delenv("SCALETORCH_B200_EP_CAPACITY", raising=False)
    assert fused_ep.default_capacity_factor() is None              # None = dropless (worst-case regions)
    monkeypatch.setenv("SCALETORCH_B200_EP_CAPACITY", "1.5")
    assert fused_ep.default_capacity_factor() == 1.5
