"""The model families compute what HuggingFace ``transformers`` computes for the same weights.

Both directions: (1) our model → ``export_hf_state_dict`` → safetensors → ``AutoModelForCausalLM.from_pretrained`` gives
the same logits and the same loss; (2) weights initialised BY ``transformers`` and saved with ``save_pretrained`` load
through our HF loader (fused q‖k‖v, gate‖up, stacked experts) and give the same logits.  This pins down the conventions
a checkpoint crosses frameworks with: tensor names, the split of the fused projections, rotate-half RoPE, per-head
QK-norm before RoPE, GQA grouping, tied embeddings, top-k routing with renormalised softmax.
"""
import os
import shutil

import pytest
import torch

transformers = pytest.importorskip("transformers")

from dist_utils import rel_err  # noqa: E402
from scaletorch_b200.trainer import ScaleTorchArguments, create_model  # noqa: E402
from scaletorch_b200.utils.checkpoint import export_hf_state_dict  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAMES = ["tiny-llama", "tiny-qwen3", "tiny-qwen3-moe"]


def _ours(model_dir, load=False, seed=0):
    args = ScaleTorchArguments(model_name_or_path=str(model_dir), micro_batch_size=1, sequence_length=16, dtype="float32",
                               use_cpu=True, load_pretrained_weights=load, seed=seed)
    torch.manual_seed(seed)
    model, mc = create_model(args, torch.float32, torch.device("cpu"))
    return model.eval(), mc


def _hf(model_dir):
    return transformers.AutoModelForCausalLM.from_pretrained(str(model_dir), dtype=torch.float32).eval()


@pytest.mark.parametrize("name", NAMES)
def test_exported_weights_give_the_same_logits_and_loss_in_transformers(tmp_path, name):
    from safetensors.torch import save_file

    d = tmp_path / name
    shutil.copytree(os.path.join(ROOT, "configs", name), d)
    ours, mc = _ours(d)
    save_file(export_hf_state_dict(ours, mc), str(d / "model.safetensors"), metadata={"format": "pt"})
    hf = _hf(d)
    ids = torch.randint(0, mc.vocab_size, (2, 25))
    x, y = ids[:, :-1], ids[:, 1:]
    with torch.no_grad():
        a, b = ours(x), hf(x).logits
    assert rel_err(a, b) < 1e-5
    ref = torch.nn.functional.cross_entropy(b.reshape(-1, b.shape[-1]), y.reshape(-1))
    assert abs(ours.loss(x, y).item() - ref.item()) < 1e-4
    if getattr(mc, "tie_word_embeddings", False):
        assert hf.get_output_embeddings().weight.data_ptr() == hf.get_input_embeddings().weight.data_ptr()


@pytest.mark.parametrize("name", NAMES)
def test_weights_initialised_by_transformers_load_through_our_loader(tmp_path, name):
    cfg = transformers.AutoConfig.from_pretrained(os.path.join(ROOT, "configs", name))
    torch.manual_seed(7)
    hf = transformers.AutoModelForCausalLM.from_config(cfg).to(torch.float32).eval()
    with torch.no_grad():                                   # make the norm weights non-trivial so a missed mapping shows
        for n, p in hf.named_parameters():
            if "norm" in n:
                p.uniform_(0.5, 1.5)
    d = tmp_path / name
    hf.save_pretrained(str(d), safe_serialization=True)
    ours, mc = _ours(d, load=True, seed=123)
    ids = torch.randint(0, mc.vocab_size, (2, 20))
    with torch.no_grad():
        assert rel_err(ours(ids), hf(ids).logits) < 1e-5


def test_gradients_agree_with_transformers(tmp_path):
    """One backward pass: d loss / d (q_proj of layer 0) from HF equals the q rows of our fused qkv gradient."""
    from safetensors.torch import save_file

    name = "tiny-qwen3"
    d = tmp_path / name
    shutil.copytree(os.path.join(ROOT, "configs", name), d)
    ours, mc = _ours(d)
    save_file(export_hf_state_dict(ours, mc), str(d / "model.safetensors"), metadata={"format": "pt"})
    hf = _hf(d)
    ids = torch.randint(0, mc.vocab_size, (2, 17))
    x, y = ids[:, :-1], ids[:, 1:]
    ours.train()
    ours.loss(x, y).backward()
    logits = hf(x).logits
    torch.nn.functional.cross_entropy(logits.reshape(-1, logits.shape[-1]), y.reshape(-1)).backward()
    hd = getattr(mc, "head_dim", None) or mc.hidden_size // mc.num_attention_heads
    nq = mc.num_attention_heads * hd

    def grad_of(p):                                           # gradients live in the flat fp32 store (main_grad)
        return p.grad if p.grad is not None else p.main_grad.view_as(p)

    g_ours = grad_of(dict(ours.named_parameters())["decoder_layers.0.attention.qkv_proj.weight"])
    g_hf = dict(hf.named_parameters())["model.layers.0.self_attn.q_proj.weight"].grad
    assert rel_err(g_ours[:nq], g_hf) < 1e-4
    g_norm = grad_of(dict(ours.named_parameters())["decoder_layers.0.attention.q_norm.weight"])
    assert rel_err(g_norm, dict(hf.named_parameters())["model.layers.0.self_attn.q_norm.weight"].grad) < 1e-4


@pytest.mark.parametrize("name", ["tiny-qwen3", "tiny-qwen3-moe"])
def test_greedy_generation_matches_transformers(tmp_path, name):
    from safetensors.torch import save_file

    d = tmp_path / name
    shutil.copytree(os.path.join(ROOT, "configs", name), d)
    ours, mc = _ours(d)
    save_file(export_hf_state_dict(ours, mc), str(d / "model.safetensors"), metadata={"format": "pt"})
    hf = _hf(d)
    prompt = torch.randint(0, mc.vocab_size, (2, 7))
    mine = ours.generate(prompt, max_new_tokens=12)
    theirs = hf.generate(prompt, max_new_tokens=12, do_sample=False, use_cache=True, pad_token_id=0, eos_token_id=None)
    assert mine.shape == (2, 19) and torch.equal(mine, theirs)
    assert not ours.training                                   # eval() models stay in eval mode
    # sampling: reproducible with a generator, stops at eos and pads with it
    g = torch.Generator().manual_seed(5)
    a = ours.generate(prompt, 6, temperature=0.8, top_k=20, generator=g)
    g = torch.Generator().manual_seed(5)
    assert torch.equal(a, ours.generate(prompt, 6, temperature=0.8, top_k=20, generator=g))
    eos = int(mine[0, 7])                                      # the first greedy token of row 0 acts as EOS
    stopped = ours.generate(prompt, 12, eos_token_id=eos)
    assert (stopped[0, 7:] == eos).all()
