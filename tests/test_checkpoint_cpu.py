"""``utils/checkpoint.py`` piece by piece (reference suite: ``tests/utils/test_checkpoint.py``): the meta-device context,
HF ↔ local name mapping for every model family, tensor-/expert-parallel slicing of HF tensors into the fused local
parameters (with a stand-in process-group manager), sharded safetensors indices, and the training checkpoint manager
(file naming per (tp, pp, ep), atomic publish, layout checks, writer election, RNG restore).
"""
import json
import os
from types import SimpleNamespace

import pytest
import torch
import torch.nn as nn

from dist_utils import TinyCfg
from scaletorch_b200.models import Llama, Qwen3, Qwen3MoE
from scaletorch_b200.parallel.context import ParallelContext
from scaletorch_b200.utils import checkpoint as C


def _fake_pgm(**kw):
    base = dict(tp_world_size=1, tp_rank=0, ep_world_size=1, ep_rank=0, pp_world_size=1, pp_rank=0, dp_rank=0, cp_rank=0)
    base.update(kw)
    return SimpleNamespace(**base)


# ------------------------------------------------------------------------------------------- meta-device context
class _WithBuffer(nn.Module):
    def __init__(self):
        super().__init__()
        self.lin = nn.Linear(4, 4)
        self.inner = nn.Sequential(nn.Linear(4, 2), nn.LayerNorm(2))
        self.register_buffer("table", torch.arange(4.0))


def test_meta_context_params_on_meta_buffers_by_request():
    with C.init_model_with_dematerialized_weights():
        m = _WithBuffer()
    assert all(p.device.type == "meta" for p in m.parameters())            # nested modules included
    assert len(list(m.parameters())) == 6
    with C.init_model_with_dematerialized_weights(include_buffers=True):
        m2 = _WithBuffer()
    assert m2.table.device.type == "meta"
    assert all(p.requires_grad for p in m2.parameters())


def test_meta_context_restores_registration_even_on_error():
    before = (nn.Module.register_parameter, nn.Module.register_buffer)
    with pytest.raises(RuntimeError, match="boom"):
        with C.init_model_with_dematerialized_weights():
            raise RuntimeError("boom")
    assert (nn.Module.register_parameter, nn.Module.register_buffer) == before
    assert nn.Linear(2, 2).weight.device.type == "cpu"                     # construction is back to normal


def test_materialise_random_init_keeps_ties_and_rebuilds_rope():
    cfg = TinyCfg(model_type="qwen3", tie_word_embeddings=True)
    with C.init_model_with_dematerialized_weights():
        m = Qwen3(cfg, ParallelContext())
    torch.manual_seed(0)
    C.init_model_with_materialized_weights(m, cfg, save_dir=None)
    assert all(p.device.type == "cpu" for p in m.parameters())
    assert m.final_proj.weight is m.embedding.weight
    assert torch.isfinite(m(torch.randint(0, cfg.vocab_size, (1, 8)))).all()
    rope = [mod for mod in m.modules() if isinstance(mod.__dict__.get("_buffers", {}).get("cos"), torch.Tensor)]
    assert rope and all(torch.isfinite(r.cos).all() and r.cos.abs().max() <= 1.0 for r in rope)


# ------------------------------------------------------------------------------------------- name mapping
def _names(model, cfg):
    mgr = C.InitializationManager(model, cfg)
    return mgr, {n: mgr.hf_names_for_local(n) for n, _ in model.named_parameters()}


def test_llama_names():
    cfg = TinyCfg(model_type="llama")
    mgr, names = _names(Llama(cfg, ParallelContext()), cfg)
    flat = {x for v in names.values() for x in v}
    assert {"model.embed_tokens.weight", "model.norm.weight", "lm_head.weight"} <= flat
    for li in range(cfg.num_hidden_layers):
        for s in ("self_attn.q_proj", "self_attn.k_proj", "self_attn.v_proj", "self_attn.o_proj", "mlp.gate_proj",
                  "mlp.up_proj", "mlp.down_proj", "input_layernorm", "post_attention_layernorm"):
            assert f"model.layers.{li}.{s}.weight" in flat
    assert not any("q_norm" in x for x in flat)                            # Llama has no QK-norm
    assert names["decoder_layers.0.attention.qkv_proj.weight"] == [f"model.layers.0.self_attn.{x}_proj.weight" for x in "qkv"]
    assert names["decoder_layers.1.mlp.gate_up_proj.weight"] == ["model.layers.1.mlp.gate_proj.weight",
                                                                 "model.layers.1.mlp.up_proj.weight"]


def test_qwen3_adds_qk_norms_and_tied_head_maps_to_embedding():
    cfg = TinyCfg(model_type="qwen3", tie_word_embeddings=True)
    m = Qwen3(cfg, ParallelContext())
    _, names = _names(m, cfg)
    flat = {x for v in names.values() for x in v}
    assert "model.layers.0.self_attn.q_norm.weight" in flat and "model.layers.1.self_attn.k_norm.weight" in flat
    assert "final_proj.weight" not in names                                # tied: one parameter, the embedding's name
    assert "lm_head.weight" not in flat


def test_moe_expert_names_follow_the_ep_rank(monkeypatch):
    cfg = TinyCfg(model_type="qwen3_moe")                                  # 4 experts
    m = Qwen3MoE(cfg, ParallelContext())
    mgr, names = _names(m, cfg)
    assert names["decoder_layers.0.moe.router.gate.weight"] == ["model.layers.0.mlp.gate.weight"]
    gu = names["decoder_layers.1.moe.experts.gate_up_proj"]
    assert gu == [f"model.layers.1.mlp.experts.{e}.{x}_proj.weight" for e in range(4) for x in ("gate", "up")]
    assert names["decoder_layers.1.moe.experts.down_proj"] == [f"model.layers.1.mlp.experts.{e}.down_proj.weight" for e in range(4)]
    monkeypatch.setattr(C, "pgm", _fake_pgm(ep_world_size=2, ep_rank=1))
    mgr2 = C.InitializationManager(m, cfg)
    assert mgr2.hf_names_for_local("decoder_layers.0.moe.experts.down_proj") == [
        f"model.layers.0.mlp.experts.{e}.down_proj.weight" for e in (2, 3)]


def test_unknown_parameter_and_pipeline_layer_ids():
    cfg = TinyCfg(model_type="llama", num_hidden_layers=4)
    m = Llama(cfg, ParallelContext())
    mgr = C.InitializationManager(m, cfg)
    with pytest.raises(KeyError):
        mgr.hf_names_for_local("decoder_layers.0.attention.rotary.inv_freq")
    with pytest.raises(KeyError):
        mgr.hf_names_for_local("something_else.weight")
    mgr.layer_ids = [2, 3]                                                 # second pipeline stage of two: local 0 -> global 2
    assert mgr.hf_names_for_local("decoder_layers.0.input_layernorm.weight") == ["model.layers.2.input_layernorm.weight"]
    assert mgr.hf_names_for_local("decoder_layers.1.mlp.down_proj.weight") == ["model.layers.3.mlp.down_proj.weight"]


# ------------------------------------------------------------------------------------------- TP / EP slicing
def test_assemble_without_tp_is_concatenation():
    cfg = TinyCfg(model_type="llama")
    mgr = C.InitializationManager(Llama(cfg, ParallelContext()), cfg)
    p = "model.layers.0."
    q, k, v = torch.randn(64, 64), torch.randn(32, 64), torch.randn(32, 64)
    hf = {p + "self_attn.q_proj.weight": q, p + "self_attn.k_proj.weight": k, p + "self_attn.v_proj.weight": v}
    assert torch.equal(mgr.assemble("decoder_layers.0.attention.qkv_proj.weight", hf), torch.cat([q, k, v]))
    emb = torch.randn(256, 64)
    assert mgr.assemble("embedding.weight", {"model.embed_tokens.weight": emb}) is not None
    assert torch.equal(mgr.assemble("embedding.weight", {"model.embed_tokens.weight": emb}), emb)


@pytest.mark.parametrize("tp_rank", [0, 1])
def test_assemble_with_tp2_slices_each_projection(monkeypatch, tp_rank):
    cfg = TinyCfg(model_type="qwen3")
    monkeypatch.setattr(C, "pgm", _fake_pgm(tp_world_size=2, tp_rank=tp_rank))
    mgr = C.InitializationManager(Qwen3(cfg, ParallelContext()), cfg)
    p = "model.layers.0."
    q, k, v = torch.randn(64, 64), torch.randn(32, 64), torch.randn(32, 64)
    hf = {p + "self_attn.q_proj.weight": q, p + "self_attn.k_proj.weight": k, p + "self_attn.v_proj.weight": v}
    got = mgr.assemble("decoder_layers.0.attention.qkv_proj.weight", hf)
    r = tp_rank
    want = torch.cat([q[32 * r:32 * (r + 1)], k[16 * r:16 * (r + 1)], v[16 * r:16 * (r + 1)]])      # [q_r; k_r; v_r]
    assert torch.equal(got, want)
    g, u = torch.randn(128, 64), torch.randn(128, 64)
    got = mgr.assemble("decoder_layers.0.mlp.gate_up_proj.weight", {p + "mlp.gate_proj.weight": g, p + "mlp.up_proj.weight": u})
    assert torch.equal(got, torch.cat([g[64 * r:64 * (r + 1)], u[64 * r:64 * (r + 1)]]))
    o = torch.randn(64, 64)
    assert torch.equal(mgr.assemble("decoder_layers.0.attention.out_proj.weight", {p + "self_attn.o_proj.weight": o}),
                       o[:, 32 * r:32 * (r + 1)])                          # row-parallel: input columns
    d = torch.randn(64, 128)
    assert torch.equal(mgr.assemble("decoder_layers.0.mlp.down_proj.weight", {p + "mlp.down_proj.weight": d}),
                       d[:, 64 * r:64 * (r + 1)])
    e = torch.randn(256, 64)
    assert torch.equal(mgr.assemble("embedding.weight", {"model.embed_tokens.weight": e}), e[128 * r:128 * (r + 1)])
    n = torch.randn(64)                                                    # norms are replicated
    assert torch.equal(mgr.assemble("decoder_layers.0.input_layernorm.weight", {p + "input_layernorm.weight": n}), n)
    qn = torch.randn(16)
    assert torch.equal(mgr.assemble("decoder_layers.0.attention.q_norm.weight", {p + "self_attn.q_norm.weight": qn}), qn)


def test_assemble_experts_are_stacked_not_tp_sliced(monkeypatch):
    cfg = TinyCfg(model_type="qwen3_moe")
    monkeypatch.setattr(C, "pgm", _fake_pgm(tp_world_size=2, tp_rank=1, ep_world_size=2, ep_rank=0))
    mgr = C.InitializationManager(Qwen3MoE(cfg, ParallelContext()), cfg)
    p = "model.layers.0.mlp.experts."
    hf = {}
    for e in range(4):
        hf[p + f"{e}.gate_proj.weight"] = torch.full((32, 64), float(e))
        hf[p + f"{e}.up_proj.weight"] = torch.full((32, 64), float(10 + e))
        hf[p + f"{e}.down_proj.weight"] = torch.full((64, 32), float(20 + e))
    gu = mgr.assemble("decoder_layers.0.moe.experts.gate_up_proj", hf)
    assert gu.shape == (2, 64, 64)                                         # this EP rank's 2 experts, full width
    assert gu[1, 0, 0] == 1.0 and gu[1, 32, 0] == 11.0                     # [gate_e; up_e]
    dn = mgr.assemble("decoder_layers.0.moe.experts.down_proj", hf)
    assert dn.shape == (2, 64, 32) and dn[0, 0, 0] == 20.0 and dn[1, 0, 0] == 21.0
    router = torch.randn(4, 64)
    assert torch.equal(mgr.assemble("decoder_layers.0.moe.router.gate.weight", {"model.layers.0.mlp.gate.weight": router}), router)


# ------------------------------------------------------------------------------------------- reading safetensors
def _export(tmp_path, cfg, model, shards: int):
    from safetensors.torch import save_file

    sd = C.export_hf_state_dict(model, cfg)
    keys = sorted(sd)
    if shards == 1:
        save_file(sd, str(tmp_path / "model.safetensors"))
        return sd
    weight_map = {}
    for i in range(shards):
        part = {k: sd[k] for k in keys[i::shards]}
        fname = f"model-{i + 1:05d}-of-{shards:05d}.safetensors"
        save_file(part, str(tmp_path / fname))
        weight_map.update({k: fname for k in part})
    json.dump({"metadata": {}, "weight_map": weight_map}, open(tmp_path / "model.safetensors.index.json", "w"))
    return sd


@pytest.mark.parametrize("shards", [1, 3])
def test_load_hf_state_from_single_file_and_sharded_index(tmp_path, shards):
    torch.manual_seed(1)
    cfg = TinyCfg(model_type="qwen3", tie_word_embeddings=True)
    src = Qwen3(cfg, ParallelContext())
    _export(tmp_path, cfg, src, shards)
    single, wmap = C._find_safetensors(str(tmp_path))
    assert (single is None) == (shards > 1) and (wmap is None) == (shards == 1)
    with C.init_model_with_dematerialized_weights():
        dst = Qwen3(cfg, ParallelContext())
    torch.manual_seed(99)
    C.init_model_with_materialized_weights(dst, cfg, save_dir=str(tmp_path), load_weights=True)
    for (n, a), (_, b) in zip(src.named_parameters(), dst.named_parameters()):
        assert torch.equal(a, b), n                                         # loaded weights are kept, not re-randomised
    assert dst.final_proj.weight is dst.embedding.weight


def test_find_safetensors_none_and_load_false_means_random_init(tmp_path):
    assert C._find_safetensors(str(tmp_path)) == (None, None)
    cfg = TinyCfg(model_type="llama")
    src = Llama(cfg, ParallelContext())
    _export(tmp_path, cfg, src, 1)
    with C.init_model_with_dematerialized_weights():
        dst = Llama(cfg, ParallelContext())
    torch.manual_seed(5)
    C.init_model_with_materialized_weights(dst, cfg, save_dir=str(tmp_path), load_weights=False)
    assert not torch.equal(dst.embedding.weight, src.embedding.weight)


def test_load_hf_state_shape_mismatch_raises(tmp_path):
    cfg = TinyCfg(model_type="llama")
    _export(tmp_path, cfg, Llama(cfg, ParallelContext()), 1)
    wide = TinyCfg(model_type="llama", intermediate_size=256)
    with C.init_model_with_dematerialized_weights():
        dst = Llama(wide, ParallelContext())
    with pytest.raises(ValueError, match="shape mismatch"):
        C.init_model_with_materialized_weights(dst, wide, save_dir=str(tmp_path), load_weights=True)


def test_partial_checkpoint_keeps_loaded_tensors_and_initialises_the_rest(tmp_path):
    from safetensors.torch import save_file

    cfg = TinyCfg(model_type="llama")
    src = Llama(cfg, ParallelContext())
    sd = C.export_hf_state_dict(src, cfg)
    sd = {k: v for k, v in sd.items() if "layers.1." not in k}             # layer 1 missing from the file
    save_file(sd, str(tmp_path / "model.safetensors"))
    with C.init_model_with_dematerialized_weights():
        dst = Llama(cfg, ParallelContext())
    torch.manual_seed(3)
    C.init_model_with_materialized_weights(dst, cfg, save_dir=str(tmp_path), load_weights=True)
    a, b = dict(src.named_parameters()), dict(dst.named_parameters())
    assert torch.equal(a["decoder_layers.0.mlp.down_proj.weight"], b["decoder_layers.0.mlp.down_proj.weight"])
    assert not torch.equal(a["decoder_layers.1.mlp.down_proj.weight"], b["decoder_layers.1.mlp.down_proj.weight"])
    assert torch.isfinite(b["decoder_layers.1.mlp.down_proj.weight"]).all()


def test_export_refuses_sharded_models(monkeypatch):
    cfg = TinyCfg(model_type="llama")
    m = Llama(cfg, ParallelContext())
    monkeypatch.setattr(C, "pgm", _fake_pgm(tp_world_size=2))
    with pytest.raises(RuntimeError, match="unsharded"):
        C.export_hf_state_dict(m, cfg)


def test_export_splits_fused_tensors_back():
    cfg = TinyCfg(model_type="qwen3_moe")
    m = Qwen3MoE(cfg, ParallelContext())
    sd = C.export_hf_state_dict(m, cfg)
    qkv = dict(m.named_parameters())["decoder_layers.0.attention.qkv_proj.weight"]
    assert torch.equal(sd["model.layers.0.self_attn.q_proj.weight"], qkv[:64])
    assert torch.equal(sd["model.layers.0.self_attn.k_proj.weight"], qkv[64:96])
    assert torch.equal(sd["model.layers.0.self_attn.v_proj.weight"], qkv[96:128])
    gu = dict(m.named_parameters())["decoder_layers.1.moe.experts.gate_up_proj"]
    assert torch.equal(sd["model.layers.1.mlp.experts.2.gate_proj.weight"], gu[2, :32])
    assert torch.equal(sd["model.layers.1.mlp.experts.2.up_proj.weight"], gu[2, 32:])
    assert all(t.is_contiguous() for t in sd.values())


# ------------------------------------------------------------------------------------------- training checkpoints
def test_checkpoint_filename_per_layout(monkeypatch):
    assert C.CheckpointManager()._filename() == "weights_tp_rank_world_size=0_1_pp_rank_world_size=0_1.pth"
    monkeypatch.setattr(C, "pgm", _fake_pgm(tp_world_size=4, tp_rank=3, pp_world_size=2, pp_rank=1))
    assert C.CheckpointManager()._filename() == "weights_tp_rank_world_size=3_4_pp_rank_world_size=1_2.pth"
    monkeypatch.setattr(C, "pgm", _fake_pgm(ep_world_size=8, ep_rank=5))
    assert C.CheckpointManager()._filename().endswith("_ep_rank_world_size=5_8.pth")       # EP ranks no longer clobber each other


def test_checkpoint_round_trip_rng_extra_and_atomic_publish(tmp_path):
    torch.manual_seed(0)
    m = nn.Sequential(nn.Linear(4, 4), nn.Linear(4, 2))
    opt = torch.optim.AdamW(m.parameters(), lr=1e-2)
    m(torch.randn(3, 4)).sum().backward()
    opt.step()
    ck = C.CheckpointManager()
    path = ck.save_checkpoint(m, opt, trained_steps=7, trained_tokens=7000, out_dir=str(tmp_path / "7"), extra={"loader": {"epoch": 1}})
    assert os.path.isfile(path) and not os.path.exists(path + ".tmp")
    expected_next = torch.rand(3)                                           # what the RNG yields right after the save
    m2 = nn.Sequential(nn.Linear(4, 4), nn.Linear(4, 2))
    opt2 = torch.optim.AdamW(m2.parameters(), lr=1e-2)
    torch.manual_seed(12345)
    steps, tokens = ck.load_checkpoint(m2, opt2, str(tmp_path / "7"))
    assert (steps, tokens) == (7, 7000) and ck.last_extra == {"loader": {"epoch": 1}}
    assert torch.equal(torch.rand(3), expected_next)                        # RNG stream continues where it was
    for a, b in zip(m.parameters(), m2.parameters()):
        assert torch.equal(a, b)
    s1, s2 = opt.state_dict()["state"], opt2.state_dict()["state"]
    assert all(torch.equal(s1[k]["exp_avg"], s2[k]["exp_avg"]) for k in s1)
    torch.manual_seed(1)
    ck.load_checkpoint(m2, None, str(tmp_path / "7"), restore_rng=False)    # optimizer optional, RNG untouched
    torch.manual_seed(1)
    a = torch.rand(1)
    torch.manual_seed(1)
    assert torch.equal(torch.rand(1), a)


def test_checkpoint_unwraps_wrappers(tmp_path):
    inner = nn.Linear(2, 2)
    wrapper = nn.Module()
    wrapper.module = inner                                                  # DataParallel-style wrapper
    ck = C.CheckpointManager()
    ck.save_checkpoint(wrapper, None, 1, 1, str(tmp_path))
    blob = torch.load(os.path.join(tmp_path, ck._filename()), weights_only=False)
    assert set(blob["model"]) == {"weight", "bias"} and blob["optimizer"] is None
    assert blob["layout"] == {"tp": 1, "pp": 1, "ep": 1}


def test_checkpoint_errors(tmp_path, monkeypatch):
    m = nn.Linear(2, 2)
    ck = C.CheckpointManager()
    with pytest.raises(FileNotFoundError, match="same TP/PP/EP layout"):
        ck.load_checkpoint(m, None, str(tmp_path))
    ck.save_checkpoint(m, None, 1, 1, str(tmp_path))
    blob = torch.load(os.path.join(tmp_path, ck._filename()), weights_only=False)
    blob["layout"] = {"tp": 2, "pp": 1, "ep": 1}
    torch.save(blob, os.path.join(tmp_path, ck._filename()))
    with pytest.raises(ValueError, match="layout"):
        ck.load_checkpoint(m, None, str(tmp_path))


@pytest.mark.parametrize("who", [dict(dp_rank=1), dict(cp_rank=1)])
def test_only_the_first_dp_cp_rank_writes(tmp_path, monkeypatch, who):
    monkeypatch.setattr(C, "pgm", _fake_pgm(**who))
    ck = C.CheckpointManager()
    assert ck.save_checkpoint(nn.Linear(2, 2), None, 1, 1, str(tmp_path / "x")) is None
    assert not os.path.exists(tmp_path / "x")


def test_to_cpu_walks_containers():
    obj = {"a": torch.ones(2), "b": [torch.zeros(1), (torch.ones(1), 3)], "c": "s"}
    out = C._to_cpu(obj)
    assert isinstance(out["b"], list) and isinstance(out["b"][1], tuple) and out["b"][1][1] == 3 and out["c"] == "s"
    assert out["a"].device.type == "cpu" and not out["a"].requires_grad


def test_async_save_snapshots_the_state_and_publishes_atomically(tmp_path):
    torch.manual_seed(0)
    m = nn.Linear(64, 64)
    opt = torch.optim.AdamW(m.parameters(), lr=1e-2)
    m(torch.randn(2, 64)).sum().backward()
    opt.step()
    want = {k: v.detach().clone() for k, v in m.state_dict().items()}
    want_m = opt.state_dict()["state"][0]["exp_avg"].clone()
    ck = C.CheckpointManager(async_save=True)
    path = ck.save_checkpoint(m, opt, 3, 300, str(tmp_path / "3"))
    with torch.no_grad():                                         # training goes on while the writer thread works
        for p in m.parameters():
            p.add_(100.0)
    opt.state_dict()["state"][0]["exp_avg"].add_(100.0)
    ck.wait()
    assert os.path.isfile(path) and not os.path.exists(path + ".tmp")
    blob = torch.load(path, weights_only=False)
    for k in want:
        assert torch.equal(blob["model"][k], want[k]), k              # the snapshot, not the later values
    assert torch.equal(blob["optimizer"]["state"][0]["exp_avg"], want_m)
    # a second save joins the first one; wait() is idempotent
    ck.save_checkpoint(m, opt, 4, 400, str(tmp_path / "4"))
    ck.wait()
    ck.wait()
    assert torch.load(os.path.join(tmp_path, "4", ck._filename()), weights_only=False)["trained_steps"] == 4


def test_async_save_reports_write_errors(tmp_path, monkeypatch):
    ck = C.CheckpointManager(async_save=True)

    def boom(*a, **k):
        raise OSError("disk full")

    monkeypatch.setattr(torch, "save", boom)
    ck.save_checkpoint(nn.Linear(2, 2), None, 1, 1, str(tmp_path / "1"))
    with pytest.raises(RuntimeError, match="disk full"):
        ck.wait()
    ck.wait()                                                     # reported once


def _fake_ckpt(root, step, shards=2, tmp=False):
    d = root / str(step)
    d.mkdir(parents=True)
    for i in range(shards):
        (d / f"weights_tp_rank_world_size={i}_2_pp_rank_world_size=0_1.pth").write_bytes(b"x")
    if tmp:
        (d / "weights_tp_rank_world_size=1_2_pp_rank_world_size=0_1.pth.tmp").write_bytes(b"x")
    return str(d)


def test_complete_checkpoint_discovery_latest_and_pruning(tmp_path):
    from scaletorch_b200.utils.checkpoint import find_complete_checkpoints, prune_checkpoints, resolve_resume_path

    assert find_complete_checkpoints(str(tmp_path / "nope")) == [] and resolve_resume_path("latest", str(tmp_path / "nope")) == ""
    assert resolve_resume_path("/some/dir", str(tmp_path)) == "/some/dir" and resolve_resume_path("", str(tmp_path)) == ""
    p2, p4, p6 = (_fake_ckpt(tmp_path, s) for s in (2, 4, 6))
    half = _fake_ckpt(tmp_path, 8, shards=1)                      # one rank had not written yet
    torn = _fake_ckpt(tmp_path, 10, tmp=True)                     # a writer was interrupted
    (tmp_path / "logs").mkdir()
    (tmp_path / "3").write_text("a file, not a checkpoint")
    assert find_complete_checkpoints(str(tmp_path)) == [(2, p2), (4, p4), (6, p6)]
    assert resolve_resume_path("latest", str(tmp_path)) == p6 and resolve_resume_path("auto", str(tmp_path)) == p6
    assert prune_checkpoints(str(tmp_path), 0) == []
    # keep the two newest complete ones; the directory being written is protected; newer incomplete ones are left alone
    assert prune_checkpoints(str(tmp_path), 2, protect=torn) == [p2]
    assert sorted(os.listdir(tmp_path)) == ["10", "3", "4", "6", "8", "logs"]
    # an incomplete directory OLDER than the newest complete checkpoint is debris
    old_torn = _fake_ckpt(tmp_path, 5, tmp=True)
    assert prune_checkpoints(str(tmp_path), 2) == [old_torn]
    assert prune_checkpoints(str(tmp_path), 1) == [p4] and find_complete_checkpoints(str(tmp_path)) == [(6, p6)]
    assert os.path.isdir(half) and os.path.isdir(torn)


def test_supervisor_and_trainer_agree_on_complete_checkpoints(tmp_path):
    import importlib.util

    from scaletorch_b200.utils.checkpoint import find_complete_checkpoints

    spec = importlib.util.spec_from_file_location("supervise", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools", "supervise.py"))
    sup = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sup)
    for s, kw in ((2, {}), (4, {"shards": 1}), (6, {}), (8, {"tmp": True})):
        _fake_ckpt(tmp_path, s, **kw)
    assert sup.complete_checkpoints(str(tmp_path)) == find_complete_checkpoints(str(tmp_path)) == [(2, str(tmp_path / "2")), (6, str(tmp_path / "6"))]


def test_inspect_checkpoint_reports_and_gates(tmp_path, capsys):
    """tools/inspect_checkpoint.py: a healthy checkpoint passes (exit 0) and is described; a missing shard, a leftover
    temporary file or a NaN weight make it fail (exit 1) with the reason in the report."""
    import importlib.util
    import shutil

    spec = importlib.util.spec_from_file_location("inspect_checkpoint", os.path.join(os.path.dirname(os.path.dirname(
        os.path.abspath(__file__))), "tools", "inspect_checkpoint.py"))
    ic = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ic)
    torch.manual_seed(0)
    m = nn.Sequential(nn.Linear(4, 4), nn.Linear(4, 2))
    opt = torch.optim.AdamW(m.parameters(), lr=1e-2)
    m(torch.randn(3, 4)).sum().backward()
    opt.step()
    good = tmp_path / "7"
    path = C.CheckpointManager().save_checkpoint(m, opt, 7, 7000, str(good), extra={"data_loader": {"batches_served": 14}})
    assert ic.main([str(good)]) == 0
    out = capsys.readouterr().out
    assert "step 7, 7000 tokens" in out and "shards 1 / 1" in out and "OK: complete and finite" in out and "'batches_served': 14" in out
    info = ic.inspect(str(good), list_tensors=True)
    assert info["total_parameters_over_shards"] == sum(p.numel() for p in m.parameters()) == 30
    assert {t["name"] for t in info["shards"][0]["tensors"]} == set(m.state_dict()) and info["shards"][0]["optimizer_bytes"] > 0
    # the same file claimed to be shard 0 of a TP2 layout: shard 1 is missing
    half = tmp_path / "8"
    half.mkdir()
    shutil.copy(path, half / "weights_tp_rank_world_size=0_2_pp_rank_world_size=0_1.pth")
    assert ic.main([str(half), "--json"]) == 1
    d = json.loads(capsys.readouterr().out)
    assert d["missing_shards"] == [[1, 0, 0]] and d["shards_expected"] == 2
    # a writer that died
    torn = tmp_path / "9"
    shutil.copytree(good, torn)
    (torn / (os.path.basename(path) + ".tmp")).write_bytes(b"x")
    assert ic.main([str(torn)]) == 1 and "TEMPORARY FILES" in capsys.readouterr().out
    # diverged weights
    with torch.no_grad():
        m[0].weight[0, 0] = float("nan")
    bad = tmp_path / "10"
    C.CheckpointManager().save_checkpoint(m, opt, 10, 10000, str(bad))
    assert ic.main([str(bad)]) == 1 and "NON-FINITE: ['0.weight']" in capsys.readouterr().out
    with pytest.raises(SystemExit):
        ic.inspect(str(tmp_path))
