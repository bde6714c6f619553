"""FSDP2 (``fully_shard``) training of the small Llama: per-block sharding, bf16 mixed precision, optional CPU offload,
explicit forward/backward prefetch and activation checkpointing (topics of the reference's
``examples/FSDP2/fsdp2_llama2_main.py:94-117,246-260`` and ``fsdp2_main.py``).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 examples/FSDP2/fsdp2_main.py --steps 5 [--cpu-offload] [--explicit-prefetch] [--ac]
"""
from __future__ import annotations

import argparse
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch.distributed.fsdp import CPUOffloadPolicy, MixedPrecisionPolicy, fully_shard

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from checkpoint import load_checkpoint, save_checkpoint  # noqa: E402
from model import ModelArgs, Transformer  # noqa: E402
from utils import describe_sharding, setup, synthetic_batch  # noqa: E402


def shard(model: Transformer, device, a) -> None:
    mp = MixedPrecisionPolicy(param_dtype=torch.bfloat16, reduce_dtype=torch.float32) if a.mixed_precision else MixedPrecisionPolicy()
    kw = dict(mp_policy=mp)
    if a.cpu_offload:
        kw["offload_policy"] = CPUOffloadPolicy()
    if a.ac:
        from torch.distributed.algorithms._checkpoint.checkpoint_wrapper import checkpoint_wrapper

        for i, blk in enumerate(model.layers):
            model.layers[i] = checkpoint_wrapper(blk)
    for blk in model.layers:                 # one communication group per block: all-gather / reduce-scatter units
        fully_shard(blk, **kw)
    fully_shard(model, **kw)                 # root group: embeddings, final norm, output head
    if a.explicit_prefetch:                  # overlap: while block i runs, gather the parameters of the next two
        blocks = list(model.layers)
        for i, blk in enumerate(blocks):
            blk.set_modules_to_forward_prefetch(blocks[i + 1:i + 3])
            blk.set_modules_to_backward_prefetch(blocks[max(i - 2, 0):i][::-1])


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--seq", type=int, default=256)
    ap.add_argument("--mixed-precision", action="store_true")
    ap.add_argument("--cpu-offload", action="store_true")
    ap.add_argument("--explicit-prefetch", action="store_true")
    ap.add_argument("--ac", action="store_true", help="activation checkpointing per block")
    ap.add_argument("--ckpt-dir", default="")
    a = ap.parse_args()
    rank, world, device = setup()
    torch.manual_seed(0)
    args = ModelArgs()
    with torch.device("meta"):               # no rank ever holds the unsharded model
        model = Transformer(args)
    shard(model, device, a)
    model.to_empty(device="cpu" if a.cpu_offload else device)
    model.init_weights()
    opt = torch.optim.AdamW(model.parameters(), lr=3e-4)
    if rank == 0:
        print(f"FSDP2 on {world} ranks ({device.type}); sharded parameters:")
    describe_sharding(model, rank)
    start = load_checkpoint(model, opt, a.ckpt_dir) if a.ckpt_dir and os.path.isdir(a.ckpt_dir) else 0
    for step in range(start, a.steps):
        x, y = synthetic_batch(a.batch, a.seq, args.vocab_size, device, seed=1000 * rank + step)
        loss = F.cross_entropy(model(x).float().flatten(0, 1), y.flatten())
        loss.backward()
        torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)
        opt.step()
        opt.zero_grad(set_to_none=True)
        if rank == 0:
            print(f"step {step:3d} loss {loss.item():.4f}", flush=True)
    if a.ckpt_dir:
        save_checkpoint(model, opt, a.ckpt_dir, a.steps)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
