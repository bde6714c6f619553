#!/usr/bin/env python
"""Analytic parameter / FLOP breakdown + timed loop → TFLOP/s and MFU (reference: ``tools/profile_mfu.py:42-148``).

``train FLOPs = 3 × forward FLOPs``; breakdown into linear layers, attention (QKᵀ and PV) and embedding / LM head.
"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaletorch_b200.trainer.config import resolve_model_config  # noqa: E402
from scaletorch_b200.utils.device import get_measured_peaks  # noqa: E402


def flop_breakdown(cfg, batch: int, seq: int) -> dict:
    H, L, V = cfg.hidden_size, cfg.num_hidden_layers, cfg.vocab_size
    hd = getattr(cfg, "head_dim", None) or H // cfg.num_attention_heads
    nq, nkv = cfg.num_attention_heads, cfg.num_key_value_heads
    T = batch * seq
    qkv = 2 * T * H * (nq + 2 * nkv) * hd
    o = 2 * T * nq * hd * H
    if getattr(cfg, "num_experts", 0):
        mlp = 2 * T * cfg.num_experts_per_tok * 3 * H * cfg.moe_intermediate_size + 2 * T * H * cfg.num_experts
    else:
        mlp = 2 * T * 3 * H * cfg.intermediate_size
    attn = 2 * 2 * batch * nq * seq * seq * hd          # QK^T + PV, full (non-causal) count like 12·L·H·d·S
    head = 2 * T * H * V
    fwd = L * (qkv + o + mlp + attn) + head
    params_linear = L * (H * (nq + 2 * nkv) * hd + nq * hd * H + 3 * H * getattr(cfg, "intermediate_size", 0))
    return {"fwd_linear": L * (qkv + o + mlp), "fwd_attention": L * attn, "fwd_lm_head": head, "fwd_total": fwd,
            "train_total": 3 * fwd, "params_linear": params_linear, "params_embed": V * H}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="Qwen3-0.6B")
    ap.add_argument("--batch_size", type=int, default=2)
    ap.add_argument("--seq_len", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--no_run", action="store_true", help="analytic numbers only")
    a = ap.parse_args()
    cfg = resolve_model_config(a.model)
    fb = flop_breakdown(cfg, a.batch_size, a.seq_len)
    for k, v in fb.items():
        print(f"{k:16s} {v / 1e12:10.3f} T")
    if a.no_run or not torch.cuda.is_available():
        return
    from bench_single import benchmark

    ms, _, tps = benchmark(a.model, a.batch_size, a.seq_len, a.steps)
    tflops = fb["train_total"] / (ms / 1e3) / 1e12
    peaks = get_measured_peaks()
    print(f"achieved {tflops:.1f} TFLOP/s = {100 * tflops / 2250:.1f}% of nominal 2.25 PF"
          + (f", {100 * tflops / peaks['bf16_tflops_sustained']:.1f}% of measured sustained" if peaks else ""))


if __name__ == "__main__":
    main()
