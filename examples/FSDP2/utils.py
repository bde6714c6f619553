"""Shared helpers of the FSDP2 examples: process-group bring-up on CUDA or CPU, synthetic token batches, inspection."""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def setup():
    """Returns ``(rank, world, device)``; NCCL on GPUs, gloo on CPU (the examples run on both)."""
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if torch.cuda.is_available():
        device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
        torch.cuda.set_device(device)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    else:
        device = torch.device("cpu")
        dist.init_process_group("gloo", rank=rank, world_size=world)
    return rank, world, device


def synthetic_batch(batch: int, seq: int, vocab: int, device, seed: int):
    g = torch.Generator(device="cpu").manual_seed(seed)
    tok = torch.randint(0, vocab, (batch, seq + 1), generator=g)
    return tok[:, :-1].to(device), tok[:, 1:].to(device)


def describe_sharding(model, rank: int) -> None:
    """Print, for a few parameters, the DTensor placements and local shapes (what FSDP2 / TP did to them)."""
    if rank != 0:
        return
    from torch.distributed.tensor import DTensor

    for i, (n, p) in enumerate(model.named_parameters()):
        if i % 9 == 0:
            if isinstance(p, DTensor):
                print(f"  {n:45s} global {tuple(p.shape)} local {tuple(p.to_local().shape)} placements {p.placements}")
            else:
                print(f"  {n:45s} {tuple(p.shape)} (not sharded)")
