"""Llama-2 under FSDP2 with a real data pipeline, LR schedule, periodic distributed checkpoints and resume
(reference: ``examples/FSDP2/fsdp2_llama2_main.py``).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 examples/FSDP2/fsdp2_llama2_main.py \
        --size tiny --steps 20 --data synthetic --ckpt-dir /tmp/llama2_ckpt --save-every 10
    # run the same command again: it resumes from step 20 (model, optimiser, scheduler, data position)
"""
from __future__ import annotations

import argparse
import math
import os
import sys
import time

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch.distributed.fsdp import MixedPrecisionPolicy, fully_shard

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from checkpoint import load_checkpoint, save_checkpoint  # noqa: E402
from data import build_loader  # noqa: E402
from llama2 import Llama2, Llama2Config  # noqa: E402
from utils import describe_sharding, setup  # noqa: E402


def cosine_with_warmup(step: int, warmup: int, total: int, floor: float = 0.1) -> float:
    if step < warmup:
        return (step + 1) / max(1, warmup)
    t = (step - warmup) / max(1, total - warmup)
    return floor + (1.0 - floor) * 0.5 * (1.0 + math.cos(math.pi * min(1.0, t)))


def build(a, device):
    cfg = Llama2Config.named(a.size)
    if a.data == "text":
        cfg.vocab_size = 257
    with torch.device("meta"):
        model = Llama2(cfg)
    mp = MixedPrecisionPolicy(param_dtype=torch.bfloat16, reduce_dtype=torch.float32) if a.bf16 else MixedPrecisionPolicy()
    for blk in model.layers:
        fully_shard(blk, mp_policy=mp, reshard_after_forward=not a.keep_params)
    fully_shard(model, mp_policy=mp)
    model.to_empty(device=device)
    torch.manual_seed(a.seed)
    model.init_weights()
    return model, cfg


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", default="tiny")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--seq", type=int, default=128)
    ap.add_argument("--lr", type=float, default=3e-4)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--grad-accum", type=int, default=1)
    ap.add_argument("--bf16", action="store_true")
    ap.add_argument("--keep-params", action="store_true", help="do not reshard after forward (ZeRO-2 style)")
    ap.add_argument("--data", default="synthetic", choices=["synthetic", "tokens", "text"])
    ap.add_argument("--path", default=None)
    ap.add_argument("--ckpt-dir", default="")
    ap.add_argument("--save-every", type=int, default=0)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    rank, world, device = setup()
    model, cfg = build(a, device)
    opt = torch.optim.AdamW(model.parameters(), lr=a.lr, betas=(0.9, 0.95), weight_decay=0.1)
    sched = torch.optim.lr_scheduler.LambdaLR(opt, lambda s: cosine_with_warmup(s, a.warmup, a.steps))
    loader = build_loader(a.data, a.path, cfg.vocab_size, a.batch, a.seq, rank, world, a.seed)
    start = 0
    if a.ckpt_dir and os.path.isdir(a.ckpt_dir):
        start = load_checkpoint(model, opt, a.ckpt_dir)
        extra = os.path.join(a.ckpt_dir, f"extra_rank{rank}.pt")
        if os.path.exists(extra):
            sd = torch.load(extra, weights_only=False)
            sched.load_state_dict(sd["sched"])
            loader.load_state_dict(sd["loader"])
    if rank == 0:
        print(f"Llama-2 {a.size}: {model.num_params() / 1e6:.1f} M parameters (global), {world} ranks, resume at step {start}")
    describe_sharding(model, rank)
    t0, tokens = time.time(), 0
    for step in range(start, a.steps):
        for micro in range(a.grad_accum):
            x, y = next(loader)
            x, y = x.to(device), y.to(device)
            # gradients are reduce-scattered only on the last micro-batch
            model.set_requires_gradient_sync(micro == a.grad_accum - 1)
            loss = F.cross_entropy(model(x).float().flatten(0, 1), y.flatten()) / a.grad_accum
            loss.backward()
            tokens += x.numel()
        norm = torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        sched.step()
        opt.zero_grad(set_to_none=True)
        if rank == 0:
            gn = norm.full_tensor().item() if hasattr(norm, "full_tensor") else float(norm)
            print(f"step {step:4d} loss {loss.item() * a.grad_accum:.4f} lr {sched.get_last_lr()[0]:.2e} grad_norm {gn:.3f} "
                  f"tok/s/rank {tokens / max(1e-9, time.time() - t0):.0f}", flush=True)
        if a.ckpt_dir and a.save_every and (step + 1) % a.save_every == 0:
            save_checkpoint(model, opt, a.ckpt_dir, step + 1)
            torch.save({"sched": sched.state_dict(), "loader": loader.state_dict()},
                       os.path.join(a.ckpt_dir, f"extra_rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
