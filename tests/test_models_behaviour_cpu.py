"""Behavioural properties of the model families on CPU (reference suites: ``tests/models/test_model_llama.py``,
``test_moe_model.py``, ``test_model.py``): causality, batch independence, GQA ≡ repeated-KV MHA, RoPE geometry, tied
embeddings, loss conventions, parameter counts, the MoE router / experts invariants and the LeNet example model.
"""
import math

import pytest
import torch
import torch.nn.functional as F

from dist_utils import TinyCfg, rel_err
from scaletorch_b200 import ops
from scaletorch_b200.models import Llama, Qwen3, Qwen3MoE
from scaletorch_b200.parallel.context import ParallelContext

FAMILIES = [(Llama, "llama"), (Qwen3, "qwen3"), (Qwen3MoE, "qwen3_moe")]


def _model(cls, mt, **kw):
    torch.manual_seed(0)
    cfg = TinyCfg(model_type=mt, **kw)
    return cls(cfg, ParallelContext()).eval(), cfg


# ------------------------------------------------------------------------------------------- dense decoder properties
@pytest.mark.parametrize("cls,mt", FAMILIES)
def test_causality_future_tokens_do_not_change_past_logits(cls, mt):
    m, cfg = _model(cls, mt)
    ids = torch.randint(0, cfg.vocab_size, (1, 24))
    other = ids.clone()
    other[0, 16:] = torch.randint(0, cfg.vocab_size, (8,))
    a, b = m(ids), m(other)
    assert rel_err(a[:, :16], b[:, :16]) < 1e-5                  # positions before the edit are untouched
    assert (a[:, 16:] - b[:, 16:]).abs().max() > 1e-4            # positions from the edit on do change


@pytest.mark.parametrize("cls,mt", [(Llama, "llama"), (Qwen3, "qwen3")])
def test_batch_rows_are_independent_and_prefix_consistent(cls, mt):
    m, cfg = _model(cls, mt)
    ids = torch.randint(0, cfg.vocab_size, (3, 20))
    full = m(ids)
    for r in range(3):
        assert rel_err(m(ids[r:r + 1]), full[r:r + 1]) < 1e-5
    assert rel_err(m(ids[:, :12]), full[:, :12]) < 1e-5          # a shorter prompt gives the same logits on its positions


def test_gqa_equals_mha_with_repeated_kv_heads():
    torch.manual_seed(0)
    B, S, Hq, Hkv, D = 2, 12, 4, 2, 16
    q, k, v = torch.randn(B, S, Hq, D), torch.randn(B, S, Hkv, D), torch.randn(B, S, Hkv, D)
    out, lse = ops.flash_attention(q, k, v, causal=True)
    kr, vr = k.repeat_interleave(Hq // Hkv, dim=2), v.repeat_interleave(Hq // Hkv, dim=2)
    ref = F.scaled_dot_product_attention(q.transpose(1, 2), kr.transpose(1, 2), vr.transpose(1, 2), is_causal=True).transpose(1, 2)
    assert rel_err(out, ref) < 1e-5
    assert lse.shape == (B, Hq, S)
    s = torch.einsum("bqhd,bkhd->bhqk", q, kr) / math.sqrt(D)
    s = s.masked_fill(torch.triu(torch.ones(S, S, dtype=torch.bool), 1), float("-inf"))
    assert rel_err(lse, torch.logsumexp(s, -1)) < 1e-5           # the log-sum-exp ring attention merges with


def test_attention_scale_and_non_causal():
    torch.manual_seed(1)
    q, k, v = (torch.randn(1, 6, 2, 8) for _ in range(3))
    out, _ = ops.flash_attention(q, k, v, causal=False, scale=0.5)
    ref = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), scale=0.5).transpose(1, 2)
    assert rel_err(out, ref) < 1e-5


def test_rope_preserves_norm_and_encodes_relative_position():
    D, S = 16, 32
    cos, sin = ops.get_cos_sin(S, D, 10000.0)
    assert cos.shape == sin.shape == (S, D) and torch.allclose(cos[0], torch.ones(D)) and torch.allclose(sin[0], torch.zeros(D))
    assert torch.allclose(cos[:, :D // 2], cos[:, D // 2:])       # halves duplicated (rotate-half convention)

    def rot(x, p):
        x1, x2 = x[..., :D // 2], x[..., D // 2:]
        return x * cos[p] + torch.cat([-x2, x1], -1) * sin[p]

    torch.manual_seed(0)
    q, k = torch.randn(D), torch.randn(D)
    for p in (0, 5, 31):
        assert torch.allclose(rot(q, p).norm(), q.norm(), atol=1e-5)
    d1 = rot(q, 7) @ rot(k, 3)
    d2 = rot(q, 20) @ rot(k, 16)                                  # same offset (4) elsewhere in the sequence
    assert torch.allclose(d1, d2, atol=1e-4)
    assert not torch.allclose(d1, rot(q, 7) @ rot(k, 5), atol=1e-4)


def test_tied_embeddings_share_storage_and_gradient():
    m, cfg = _model(Qwen3, "qwen3", tie_word_embeddings=True)
    assert m.final_proj.weight is m.embedding.weight
    n_tied = sum(p.numel() for p in m.parameters())
    m2, _ = _model(Qwen3, "qwen3", tie_word_embeddings=False)
    assert sum(p.numel() for p in m2.parameters()) - n_tied == cfg.vocab_size * cfg.hidden_size
    ids = torch.randint(0, cfg.vocab_size, (1, 9))
    m.train()
    m.loss(ids[:, :-1], ids[:, 1:]).backward()
    g = m.embedding.weight.grad
    assert g is not None and g.abs().sum() > 0
    used = ids[0, :-1].unique()
    unused = torch.tensor([i for i in range(cfg.vocab_size) if i not in set(used.tolist())][:5])
    assert g[unused].abs().sum() > 0                              # rows never looked up still get the LM-head gradient


@pytest.mark.parametrize("cls,mt", FAMILIES)
def test_loss_conventions(cls, mt):
    m, cfg = _model(cls, mt)
    ids = torch.randint(0, cfg.vocab_size, (2, 13))
    x, y = ids[:, :-1], ids[:, 1:].clone()
    base = m.loss(x, y)
    assert 0.5 * math.log(cfg.vocab_size) < base.item() < 2.0 * math.log(cfg.vocab_size)     # random init ≈ ln V
    assert torch.allclose(m.loss(x, y, loss_scale=0.25), base * 0.25, rtol=1e-5)
    y_ign = y.clone()
    y_ign[:, :6] = -100                                           # ignored positions do not count, mean over the rest
    logits = m(x)
    ref = F.cross_entropy(logits.reshape(-1, cfg.vocab_size), y_ign.reshape(-1), ignore_index=-100)
    assert torch.allclose(m.loss(x, y_ign), ref, rtol=1e-4)


def test_head_dim_override_and_kv_head_validation():
    m, cfg = _model(Qwen3, "qwen3", head_dim=32)                  # Qwen3: head_dim independent of hidden / heads
    qkv = dict(m.named_parameters())["decoder_layers.0.attention.qkv_proj.weight"]
    assert qkv.shape == ((4 + 2 * 2) * 32, cfg.hidden_size)
    with pytest.raises((ValueError, AssertionError)):
        Llama(TinyCfg(model_type="llama", num_attention_heads=4, num_key_value_heads=3), ParallelContext())


def test_gradients_reach_every_parameter():
    for cls, mt in FAMILIES:
        m, cfg = _model(cls, mt)
        m.train()
        ids = torch.randint(0, cfg.vocab_size, (2, 17))
        loss = m.loss(ids[:, :-1], ids[:, 1:])
        aux = m.get_aux_loss()
        (loss + (aux if aux is not None else 0.0)).backward()
        dead = [n for n, p in m.named_parameters() if p.grad is None or not torch.isfinite(p.grad).all()]
        assert not dead, (mt, dead)


# ------------------------------------------------------------------------------------------- MoE invariants
def test_router_reference_invariants():
    from scaletorch_b200.ops.moe_routing import route_ref

    torch.manual_seed(0)
    T, E, k = 64, 8, 2
    logits = torch.randn(T, E)
    w, idx, pos, counts, aux = route_ref(logits, k, aux_coef=0.01, training=True)
    assert w.shape == idx.shape == pos.shape == (T, k) and counts.shape == (E,)
    assert torch.allclose(w.sum(-1), torch.ones(T), atol=1e-6) and (w > 0).all()
    assert (idx[:, 0] != idx[:, 1]).all()                          # k distinct experts per token
    assert torch.equal(idx[:, 0], logits.argmax(-1))               # first choice is the arg-max, weights descending
    assert (w[:, 0] >= w[:, 1]).all()
    assert int(counts.sum()) == T * k and torch.equal(counts, torch.bincount(idx.reshape(-1), minlength=E))
    for e in range(E):                                             # positions: a bijection onto 0..count-1 per expert
        p = pos.reshape(-1)[idx.reshape(-1) == e]
        assert sorted(p.tolist()) == list(range(int(counts[e])))
    assert aux.item() > 0
    _, _, _, _, aux0 = route_ref(logits, k, aux_coef=0.0, training=True)
    _, _, _, _, aux_eval = route_ref(logits, k, aux_coef=0.01, training=False)
    assert aux0.item() == 0.0 and aux_eval.item() == 0.0


def test_aux_loss_is_minimal_for_balanced_routing():
    from scaletorch_b200.ops.moe_routing import route_ref

    E, k, T = 4, 1, 400
    balanced = torch.zeros(T, E)
    balanced[torch.arange(T), torch.arange(T) % E] = 5.0           # every expert gets T/E tokens, confidently
    collapsed = torch.zeros(T, E)
    collapsed[:, 0] = 5.0                                          # everything to expert 0
    a_bal = route_ref(balanced, k, 1.0, True)[4].item()
    a_col = route_ref(collapsed, k, 1.0, True)[4].item()
    assert a_bal == pytest.approx(1.0, abs=0.05)                   # Switch loss: E · Σ f_e p_e = 1 when uniform
    assert a_col > 2.5 * a_bal


def test_moe_layer_equals_explicit_per_token_mixture():
    from scaletorch_b200.models.model_qwen3_moe import MoELayer

    torch.manual_seed(0)
    cfg = TinyCfg(model_type="qwen3_moe")
    layer = MoELayer(cfg, ParallelContext()).eval()
    x = torch.randn(10, cfg.hidden_size)
    y = layer(x)
    w, idx, _ = layer.router(x)
    want = torch.zeros_like(x)
    I = cfg.moe_intermediate_size
    for t in range(x.shape[0]):
        for j in range(idx.shape[1]):
            e = int(idx[t, j])
            gu = layer.experts.gate_up_proj[e] @ x[t]
            h = F.silu(gu[:I]) * gu[I:]
            want[t] += w[t, j] * (layer.experts.down_proj[e] @ h)
    assert rel_err(y, want) < 1e-5


def test_moe_output_is_permutation_equivariant_over_tokens():
    from scaletorch_b200.models.model_qwen3_moe import MoELayer

    torch.manual_seed(1)
    cfg = TinyCfg(model_type="qwen3_moe")
    layer = MoELayer(cfg, ParallelContext()).eval()
    x = torch.randn(16, cfg.hidden_size)
    perm = torch.randperm(16)
    assert rel_err(layer(x[perm]), layer(x)[perm]) < 1e-5         # the sort-by-expert dispatch is invisible outside


def test_moe_rejects_indivisible_expert_parallelism():
    from scaletorch_b200.models.model_qwen3_moe import MoELayer

    ctx = ParallelContext()
    ctx.ep_size = 3
    with pytest.raises(ValueError, match="divisible"):
        MoELayer(TinyCfg(model_type="qwen3_moe"), ctx)


def test_moe_model_sparse_layers_and_param_count():
    m, cfg = _model(Qwen3MoE, "qwen3_moe")
    names = [n for n, _ in m.named_parameters()]
    assert sum("moe.experts.gate_up_proj" in n for n in names) == cfg.num_hidden_layers
    assert not any("mlp.gate_up_proj" in n for n in names)         # every layer sparse at decoder_sparse_step = 1
    experts = sum(p.numel() for n, p in m.named_parameters() if ".experts." in n)
    assert experts == cfg.num_hidden_layers * cfg.num_experts * 3 * cfg.hidden_size * cfg.moe_intermediate_size


# ------------------------------------------------------------------------------------------- small models
def test_lenet_shapes_and_training_step():
    from scaletorch_b200.models.lenet import LeNet

    torch.manual_seed(0)
    net = LeNet()
    x = torch.randn(4, 1, 28, 28)
    out = net(x)
    assert out.shape == (4, 10)
    opt = torch.optim.SGD(net.parameters(), lr=0.1)
    y = torch.tensor([1, 2, 3, 4])
    l0 = F.cross_entropy(net(x), y)
    for _ in range(20):
        opt.zero_grad()
        F.cross_entropy(net(x), y).backward()
        opt.step()
    assert F.cross_entropy(net(x), y) < l0
    assert LeNet(num_classes=7)(x).shape == (4, 7)


def test_rmsnorm_module_matches_formula():
    from scaletorch_b200.models.attention_utils import RMSNorm

    torch.manual_seed(0)
    n = RMSNorm(32, eps=1e-5)
    with torch.no_grad():
        n.weight.uniform_(0.5, 1.5)
    x = torch.randn(3, 7, 32)
    want = x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-5) * n.weight
    assert rel_err(n(x), want) < 1e-6
    assert "32" in n.extra_repr()
