"""HF transformers Qwen3 / Qwen3-MoE with the framework's ops patched in must agree with the stock modules."""
import pytest
import torch

transformers = pytest.importorskip("transformers")


def _tiny(moe: bool):
    common = dict(vocab_size=128, hidden_size=64, num_hidden_layers=2, num_attention_heads=2, num_key_value_heads=1,
                  head_dim=32, intermediate_size=128, max_position_embeddings=64, rms_norm_eps=1e-6, tie_word_embeddings=False)
    if moe:
        from transformers import Qwen3MoeConfig, Qwen3MoeForCausalLM

        cfg = Qwen3MoeConfig(num_experts=4, num_experts_per_tok=2, moe_intermediate_size=32, norm_topk_prob=True,
                             decoder_sparse_step=1, mlp_only_layers=[], **common)
        return Qwen3MoeForCausalLM(cfg)
    from transformers import Qwen3Config, Qwen3ForCausalLM

    return Qwen3ForCausalLM(Qwen3Config(**common))


@pytest.mark.parametrize("moe", [False, True])
def test_patched_hf_model_matches_stock(moe):
    from scaletorch_b200.models.hf_patch import apply_b200_patch, remove_b200_patch

    torch.manual_seed(0)
    model = _tiny(moe).eval()
    for p in model.parameters():            # HF leaves expert tensors uninitialised (torch.empty) in tiny random models
        if not torch.isfinite(p).all() or p.abs().max() > 10:
            torch.nn.init.normal_(p, std=0.02)
    ids = torch.randint(0, 128, (2, 16))
    ref = model(input_ids=ids, labels=ids)
    ref.loss.backward()
    g_ref = {n: p.grad.clone() for n, p in model.named_parameters() if p.grad is not None}
    model.zero_grad()
    names = apply_b200_patch()
    try:
        assert "Qwen3RMSNorm" in names and ("Qwen3MoeExperts" in names or not moe)
        out = model(input_ids=ids, labels=ids)
        out.loss.backward()
    finally:
        remove_b200_patch()
    assert torch.allclose(out.logits, ref.logits, atol=2e-4, rtol=1e-3), (out.logits - ref.logits).abs().max()
    assert abs(out.loss.item() - ref.loss.item()) < 1e-4
    for n, p in model.named_parameters():
        if n in g_ref:
            assert torch.allclose(p.grad, g_ref[n], atol=2e-4, rtol=1e-2), n
    # the patch is gone again
    again = model(input_ids=ids)
    assert torch.allclose(again.logits, ref.logits, atol=1e-6)
