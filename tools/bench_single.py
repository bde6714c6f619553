#!/usr/bin/env python
"""Single-GPU speed / memory benchmark (reference: ``tools/bench_single.py:30-135``, there hard-coded to ``npu``).

Meta-init + materialise, bf16, synthetic ``torch.randint`` tokens, fused linear+CE loss, global-norm clip,
fused AdamW; 3 warm-up steps, N timed steps with CUDA events; prints step ms, tokens/s, peak / reserved memory.
``--cpu`` is a dry run of the same loop on the host (fp32, wall clock, no memory figures): it checks the plumbing, not speed.
"""
import argparse
import gc
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaletorch_b200.trainer import ScaleTorchArguments, create_model, create_optimizer  # noqa: E402
from scaletorch_b200.utils.misc import get_mfu, get_num_params  # noqa: E402


def benchmark(model_path="Qwen3-0.6B", batch_size=4, seq_len=2048, steps=10, warmup=3, gradient_checkpointing=False,
              device_id=0, cpu=False):
    if cpu:
        return _benchmark_cpu(model_path, batch_size, seq_len, steps, warmup, gradient_checkpointing)
    if not torch.cuda.is_available():
        raise SystemExit("bench_single.py: no CUDA device (pass --cpu for a host dry run)")
    device = torch.device("cuda", device_id)
    torch.cuda.set_device(device)
    torch.cuda.empty_cache()
    gc.collect()
    cfg = ScaleTorchArguments(model_name_or_path=model_path, micro_batch_size=batch_size, sequence_length=seq_len,
                              synthetic_data=True, learning_rate=3e-4, max_grad_norm=1.0)
    model, mc = create_model(cfg, torch.bfloat16, device)
    model.train()
    n = get_num_params(model)
    print(f"Model params: {n / 1e6:.1f}M")
    opt = create_optimizer(model, cfg, device)

    def train_step():
        x = torch.randint(0, mc.vocab_size, (batch_size, seq_len + 1), device=device)
        opt.zero_grad()
        loss = model.loss(x[:, :-1], x[:, 1:], gradient_checkpointing=gradient_checkpointing)
        loss.backward()
        opt.step()
        return loss

    for _ in range(warmup):
        train_step()
    torch.cuda.synchronize()
    torch.cuda.reset_peak_memory_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    losses = [train_step() for _ in range(steps)]
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    tps = batch_size * seq_len / (ms / 1e3)
    print(f"Config: BS={batch_size}, SEQ={seq_len}, GC={gradient_checkpointing}")
    print(f"Step time: {ms:.1f}ms")
    print(f"Tokens/s: {tps:.0f}")
    print(f"MFU (nominal 2.25 PF): {get_mfu(tps, n, mc, 2.25e15, seq_len):.1f}%")
    print(f"Peak Memory: {torch.cuda.max_memory_allocated() / 1e9:.2f} GB")
    print(f"Reserved Memory: {torch.cuda.memory_reserved() / 1e9:.2f} GB")
    print(f"Avg Loss: {sum(float(l) for l in losses) / len(losses):.4f}")
    return ms, torch.cuda.max_memory_allocated() / 1e9, tps


def _benchmark_cpu(model_path, batch_size, seq_len, steps, warmup, gradient_checkpointing):
    device = torch.device("cpu")
    cfg = ScaleTorchArguments(model_name_or_path=model_path, micro_batch_size=batch_size, sequence_length=seq_len,
                              synthetic_data=True, learning_rate=3e-4, max_grad_norm=1.0, use_cpu=True)
    model, mc = create_model(cfg, torch.float32, device)
    model.train()
    n = get_num_params(model)
    print(f"Model params: {n / 1e6:.1f}M")
    opt = create_optimizer(model, cfg, device)
    losses = []
    for i in range(warmup + steps):
        if i == warmup:
            t0 = time.perf_counter()
        x = torch.randint(0, mc.vocab_size, (batch_size, seq_len + 1))
        opt.zero_grad()
        loss = model.loss(x[:, :-1], x[:, 1:], gradient_checkpointing=gradient_checkpointing)
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
    ms = (time.perf_counter() - t0) * 1e3 / steps
    tps = batch_size * seq_len / (ms / 1e3)
    print(f"Config: BS={batch_size}, SEQ={seq_len}, GC={gradient_checkpointing} (CPU dry run, fp32)")
    print(f"Step time: {ms:.1f}ms")
    print(f"Tokens/s: {tps:.0f}")
    print(f"Avg Loss: {sum(losses[warmup:]) / steps:.4f}")
    return ms, 0.0, tps


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--model_path", default="Qwen3-0.6B")
    ap.add_argument("--batch_size", type=int, default=4)
    ap.add_argument("--seq_len", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--gradient_checkpointing", action="store_true")
    ap.add_argument("--device_id", type=int, default=0)
    ap.add_argument("--cpu", action="store_true", help="host dry run (fp32, wall clock)")
    a = ap.parse_args()
    benchmark(a.model_path, a.batch_size, a.seq_len, a.steps, gradient_checkpointing=a.gradient_checkpointing,
              device_id=a.device_id, cpu=a.cpu)
