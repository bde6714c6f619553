"""Timing script for the nanoGPT-style GPT-MoE (``scaletorch_b200/models/moe.py``): dense vs MoE at equal width, forward
and forward+backward, tokens/s and the router's expert-usage statistics (reference: ``tests/models/benchmark_moe.py``,
a timing script that lives in its test tree).

    python tools/benchmark_moe.py [--device cuda] [--n_embd 512] [--n_layer 4] [--n_exp 8] [--top_k 2] [--batch 8] [--block 256]
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaletorch_b200.models.moe import GPT, GPTConfig, analyze_moe_usage  # noqa: E402


def timeit(fn, iters: int, device: torch.device) -> float:
    for _ in range(2):
        fn()
    if device.type == "cuda":
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters
    t0 = time.perf_counter()
    for _ in range(iters):
        fn()
    return (time.perf_counter() - t0) * 1e3 / iters


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    ap.add_argument("--n_embd", type=int, default=256)
    ap.add_argument("--n_layer", type=int, default=4)
    ap.add_argument("--n_head", type=int, default=4)
    ap.add_argument("--n_exp", type=int, default=8)
    ap.add_argument("--top_k", type=int, default=2)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--block", type=int, default=128)
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    dev = torch.device(a.device)
    idx = torch.randint(0, 1024, (a.batch, a.block), device=dev)
    rows = []
    for name, n_exp in (("dense", 1), (f"moe{a.n_exp}x top{a.top_k}", a.n_exp)):
        cfg = GPTConfig(block_size=a.block, vocab_size=1024, n_layer=a.n_layer, n_head=a.n_head, n_embd=a.n_embd,
                        use_moe=n_exp > 1, n_experts=max(n_exp, 1), top_k=a.top_k if n_exp > 1 else 1)
        model = GPT(cfg).to(dev)

        def fwd():
            with torch.no_grad():
                model(idx, idx)

        def fwdbwd():
            model.zero_grad(set_to_none=True)
            _, loss = model(idx, idx)
            loss.backward()

        f, fb = timeit(fwd, a.iters, dev), timeit(fwdbwd, a.iters, dev)
        rows.append((name, model.get_num_params() / 1e6, f, fb, a.batch * a.block / fb * 1e3))
        if n_exp > 1:
            usage = analyze_moe_usage(model)
            print(f"{name}: {usage}")
    print(f"{'model':24s} {'params (M)':>10s} {'fwd ms':>10s} {'fwd+bwd ms':>12s} {'tok/s':>12s}")
    for r in rows:
        print(f"{r[0]:24s} {r[1]:10.2f} {r[2]:10.2f} {r[3]:12.2f} {r[4]:12.0f}")


if __name__ == "__main__":
    main()
