"""Model families, attention variants, GPT-MoE, checkpoint / HF layout round trips (CPU)."""
import os

import pytest
import torch

from dist_utils import TinyCfg, rel_err
from scaletorch_b200.models import Llama, Qwen3, Qwen3MoE
from scaletorch_b200.parallel.context import ParallelContext


@pytest.mark.parametrize("cls,mt", [(Llama, "llama"), (Qwen3, "qwen3"), (Qwen3MoE, "qwen3_moe")])
def test_forward_loss_backward_and_checkpointing(cls, mt):
    torch.manual_seed(0)
    cfg = TinyCfg(model_type=mt, tie_word_embeddings=(mt == "qwen3"))
    m = cls(cfg, ParallelContext())
    ids = torch.randint(0, cfg.vocab_size, (2, 17))
    x, y = ids[:, :-1], ids[:, 1:]
    logits = m(x)
    assert logits.shape == (2, 16, cfg.vocab_size)
    ref = torch.nn.functional.cross_entropy(logits.reshape(-1, cfg.vocab_size), y.reshape(-1))
    loss = m.loss(x, y)
    assert abs(loss.item() - ref.item()) < 1e-5
    loss.backward()
    g1 = {n: p.grad.clone() for n, p in m.named_parameters()}
    m.zero_grad()
    m.loss(x, y, gradient_checkpointing=True).backward()
    for n, p in m.named_parameters():
        assert rel_err(p.grad, g1[n]) < 1e-5, n
    if mt == "qwen3":
        assert m.final_proj.weight is m.embedding.weight
    if mt == "qwen3_moe":
        m.train()
        m.loss(x, y)
        assert m.get_aux_loss() is not None and m.get_aux_loss().item() > 0
    else:
        assert m.get_aux_loss() is None


def test_head_divisibility_is_validated():
    with pytest.raises(ValueError):
        Qwen3(TinyCfg(model_type="qwen3", num_attention_heads=4, num_key_value_heads=2), ParallelContext(tp_size=4))


def test_rope_helpers():
    from scaletorch_b200.models.attention_utils import apply_rotary_pos_emb, get_cos_sin

    cos, sin = get_cos_sin(16, 8, 10000.0)
    assert cos.shape == (16, 8) and cos.dtype == torch.float32
    assert torch.allclose(cos[:, :4], cos[:, 4:]) and torch.allclose(cos[0], torch.ones(8))
    x = torch.randn(2, 3, 16, 8)
    y = apply_rotary_pos_emb(x, cos, sin)
    assert torch.allclose(y.norm(dim=-1), x.norm(dim=-1), atol=1e-5)      # rotation preserves norms
    assert torch.allclose(y[:, :, 0], x[:, :, 0], atol=1e-6)              # position 0 is the identity


def test_attention_backend_registry():
    from scaletorch_b200.models.attention_utils import (get_attention_backend, list_attention_backends,
                                                        register_attention_backend, resolve_attention_backend_name)

    assert {"ring", "flash", "sdpa"} <= set(list_attention_backends())
    assert resolve_attention_backend_name(True, True) == "ring"
    assert resolve_attention_backend_name(False, True) == "flash" and resolve_attention_backend_name(False, False) == "sdpa"
    register_attention_backend("mine", lambda q, k, v, causal=True, scale=None: q)
    assert get_attention_backend("mine") is not None
    with pytest.raises(KeyError):
        get_attention_backend("nope")
    q, k, v = torch.randn(1, 8, 4, 16), torch.randn(1, 8, 2, 16), torch.randn(1, 8, 2, 16)
    assert rel_err(get_attention_backend("flash")(q, k, v), get_attention_backend("sdpa")(q, k, v)) < 1e-5


def test_standalone_attention_variants():
    from scaletorch_b200.models.attention import (GroupQueryAttention, MultiHeadAttention, MultiHeadLatentAttention,
                                                  MultiQueryAttention)

    x = torch.randn(2, 10, 64)
    mask = torch.ones(2, 10)
    mask[:, 7:] = 0
    for m in (MultiHeadAttention(64, 4), MultiQueryAttention(64, 4, dropout=0.0), GroupQueryAttention(64, 4, 2, dropout=0.0),
              MultiHeadLatentAttention(64, 4, 16, 16)):
        y = m(x)
        y2, w = m(x, return_attention_weights=True)
        assert y.shape == x.shape and w.shape == (2, 4, 10, 10)
        assert torch.allclose(y, y2, atol=1e-5) and torch.allclose(w.sum(-1), torch.ones(2, 4, 10), atol=1e-5)
        _, wm = m(x, attention_mask=mask, return_attention_weights=True)
        assert wm[..., 7:].abs().max() == 0
        y.sum().backward()
    with pytest.raises(ValueError):
        MultiHeadAttention(65, 4)
    with pytest.raises(ValueError):
        GroupQueryAttention(64, 4, 3)
    with pytest.raises(ValueError):
        MultiHeadAttention(64, 4)(torch.randn(2, 64))


def test_gpt_moe():
    from scaletorch_b200.models.moe import GPT, GPTConfig, analyze_moe_usage, get_moe_layer_info

    with pytest.raises(ValueError):
        GPTConfig(n_embd=30, n_head=4)
    with pytest.raises(ValueError):
        GPTConfig(use_moe=True, n_experts=2, top_k=3)
    cfg = GPTConfig(block_size=32, vocab_size=128, n_layer=2, n_head=2, n_embd=32, use_moe=True, n_experts=4, top_k=2,
                    use_aux_loss=True, use_router_z_loss=True)
    m = GPT(cfg)
    x = torch.randint(0, 128, (2, 16))
    logits, loss = m(x, x)
    assert logits.shape == (2, 16, 128) and loss.item() > 0
    loss.backward()
    assert analyze_moe_usage(m)["moe_layers"] == [1] and get_moe_layer_info(m, 1)["n_experts"] == 4
    assert get_moe_layer_info(m, 0) is None
    m.eval()
    assert m.generate(x[:, :4], 5, top_k=10).shape == (2, 9)
    last, none = m(x)
    assert last.shape == (2, 1, 128) and none is None
    m.crop_block_size(16)
    with pytest.raises(ValueError):
        m(torch.randint(0, 128, (1, 17)))
    assert 0 < m.estimate_mfu(1, 1.0, peak_flops=1e12) < 1


def test_hf_layout_export_load_roundtrip_keeps_weights(tmp_path):
    """export_hf_state_dict → safetensors → create_model(load_pretrained_weights) reproduces the same logits
    (the reference re-randomises after loading)."""
    import shutil

    from safetensors.torch import save_file

    from scaletorch_b200.trainer import ScaleTorchArguments, create_model
    from scaletorch_b200.utils.checkpoint import export_hf_state_dict

    for name in ("tiny-qwen3", "tiny-qwen3-moe", "tiny-llama"):
        src = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "configs", name)
        d = tmp_path / name
        shutil.copytree(src, d)
        cfg = ScaleTorchArguments(model_name_or_path=str(d), micro_batch_size=1, sequence_length=16, dtype="float32", use_cpu=True)
        m1, mc = create_model(cfg, torch.float32, torch.device("cpu"))
        sd = export_hf_state_dict(m1, mc)
        assert "model.layers.0.self_attn.q_proj.weight" in sd and "model.embed_tokens.weight" in sd
        if name == "tiny-qwen3-moe":
            assert "model.layers.1.mlp.experts.3.down_proj.weight" in sd and "model.layers.0.mlp.gate.weight" in sd
        if getattr(mc, "tie_word_embeddings", False):
            assert "lm_head.weight" not in sd
        save_file(sd, str(d / "model.safetensors"))
        cfg2 = ScaleTorchArguments(model_name_or_path=str(d), micro_batch_size=1, sequence_length=16, dtype="float32",
                                   use_cpu=True, load_pretrained_weights=True, seed=999)
        torch.manual_seed(999)
        m2, _ = create_model(cfg2, torch.float32, torch.device("cpu"))
        ids = torch.randint(0, mc.vocab_size, (1, 16))
        assert rel_err(m2(ids), m1(ids)) < 1e-6, name


def test_training_checkpoint_save_resume(tmp_path):
    from scaletorch_b200.trainer import ScaleTorchArguments, create_model, create_optimizer
    from scaletorch_b200.utils.checkpoint import CheckpointManager

    cfg = ScaleTorchArguments(model_name_or_path="tiny-qwen3", micro_batch_size=1, sequence_length=16, dtype="float32", use_cpu=True)
    dev = torch.device("cpu")
    m, mc = create_model(cfg, torch.float32, dev)
    opt = create_optimizer(m, cfg, dev)
    ids = torch.randint(0, mc.vocab_size, (1, 17))
    for _ in range(2):
        opt.zero_grad()
        m.loss(ids[:, :-1], ids[:, 1:]).backward()
        opt.step()
    ck = CheckpointManager()
    path = ck.save_checkpoint(m, opt, 2, 32, str(tmp_path / "2"), extra={"x": 1})
    assert os.path.basename(path) == "weights_tp_rank_world_size=0_1_pp_rank_world_size=0_1.pth"
    ref_logits = m(ids[:, :-1]).detach().clone()
    m2, _ = create_model(ScaleTorchArguments(model_name_or_path="tiny-qwen3", micro_batch_size=1, sequence_length=16,
                                             dtype="float32", use_cpu=True, seed=5), torch.float32, dev)
    opt2 = create_optimizer(m2, cfg, dev)
    steps, toks = ck.load_checkpoint(m2, opt2, str(tmp_path / "2"))
    assert (steps, toks) == (2, 32) and ck.last_extra == {"x": 1}
    assert rel_err(m2(ids[:, :-1]), ref_logits) < 1e-6
    assert opt2.step_count == 2 and torch.equal(opt2.exp_avg, opt.exp_avg)
    # both continue identically
    for mm, oo in ((m, opt), (m2, opt2)):
        oo.zero_grad()
        mm.loss(ids[:, :-1], ids[:, 1:]).backward()
        oo.step()
    assert rel_err(m2(ids[:, :-1]), m(ids[:, :-1])) < 1e-6
    with pytest.raises(FileNotFoundError):
        ck.load_checkpoint(m2, opt2, str(tmp_path / "nope"))


def test_meta_device_context_allocates_nothing():
    from scaletorch_b200.utils.checkpoint import init_model_with_dematerialized_weights
    from scaletorch_b200.utils.misc import assert_no_meta_tensors

    with init_model_with_dematerialized_weights():
        m = Qwen3(TinyCfg(model_type="qwen3"), ParallelContext())
    assert all(p.device.type == "meta" for p in m.parameters())
    with pytest.raises(RuntimeError):
        assert_no_meta_tensors(m)


def test_bucket_manager_layout():
    from scaletorch_b200.parallel.data_parallel.bucket import BucketManager
    from scaletorch_b200.runtime.flat_params import FlatParamStore

    m = torch.nn.Sequential(*[torch.nn.Linear(32, 32, bias=False) for _ in range(5)])
    store = FlatParamStore(m)
    bm = BucketManager(store, None, None, bucket_size=2500)
    assert len(bm.buckets) == 3 and sum(len(b.params) for b in bm.buckets) == 5
    assert bm.buckets[0].params[0] is m[4].weight          # reverse registration order ≈ backward order
    for b in bm.buckets:
        assert b.grad_data.data_ptr() >= store.grad_flat.data_ptr() and b.numel >= sum(p.numel() for p in b.params)
    p = m[4].weight
    bm.mark_param_as_ready(p, sync=False)
    with pytest.raises(RuntimeError):
        bm.mark_param_as_ready(p, sync=False)
    bm.reset()
    bm.mark_param_as_ready(p, sync=False)
