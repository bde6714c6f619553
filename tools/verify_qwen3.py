#!/usr/bin/env python
"""Checks a local Qwen3 checkpoint / config end to end (reference: ``tools/verify_qwen3.py:19-97``): config print,
meta-init, materialise (loading *and keeping* HF weights when present), tied-weight ``data_ptr`` check, finite forward,
forward/backward gradient check."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaletorch_b200.trainer import ScaleTorchArguments, create_model  # noqa: E402


def verify(path: str, device: torch.device, dtype: torch.dtype) -> bool:
    cfg = ScaleTorchArguments(model_name_or_path=path, micro_batch_size=1, sequence_length=128, synthetic_data=True,
                              load_pretrained_weights=True, dtype="bfloat16" if dtype == torch.bfloat16 else "float32",
                              use_cpu=device.type == "cpu")
    model, mc = create_model(cfg, dtype, device)
    print(f"[{path}] type={mc.model_type} layers={mc.num_hidden_layers} hidden={mc.hidden_size} heads={mc.num_attention_heads}"
          f"/{mc.num_key_value_heads} vocab={mc.vocab_size} tie={getattr(mc, 'tie_word_embeddings', False)}")
    ok = True
    if getattr(mc, "tie_word_embeddings", False):
        tied = model.embedding.weight.data_ptr() == model.final_proj.weight.data_ptr()
        print(f"  tied embedding/lm_head share storage: {tied}")
        ok &= tied
    ids = torch.randint(0, mc.vocab_size, (1, 129), device=device)
    with torch.no_grad():
        logits = model(ids[:, :-1])
    fin = bool(torch.isfinite(logits.float()).all())
    print(f"  forward finite: {fin}  logits {tuple(logits.shape)}")
    ok &= fin
    loss = model.loss(ids[:, :-1], ids[:, 1:])
    loss.backward()
    store = getattr(model, "_stb_store", None)
    g = store.grad_flat if store is not None else torch.cat([p.grad.reshape(-1).float() for p in model.parameters() if p.grad is not None])
    gfin = bool(torch.isfinite(g).all()) and float(g.abs().sum()) > 0
    print(f"  loss {float(loss.detach()):.4f}  gradients finite and non-zero: {gfin}")
    return ok and gfin


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("paths", nargs="*", default=["tiny-qwen3"])
    ap.add_argument("--cpu", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cpu") if a.cpu or not torch.cuda.is_available() else torch.device("cuda", 0)
    dt = torch.float32 if dev.type == "cpu" else torch.bfloat16
    res = [verify(p, dev, dt) for p in a.paths]
    sys.exit(0 if all(res) else 1)
