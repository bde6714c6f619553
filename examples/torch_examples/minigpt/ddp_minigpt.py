"""Character-level miniGPT under DDP with *snapshot resume*: every ``--save-every`` steps rank 0 writes model + optimiser
+ step; a relaunched job (torchrun restart, pre-emption) picks up from the snapshot (reference:
``examples/torch_examples/minigpt``).  The text is generated locally (no downloads).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 examples/torch_examples/minigpt/ddp_minigpt.py --steps 40 --snapshot /tmp/minigpt.pt
"""
import argparse
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F
from torch.nn.parallel import DistributedDataParallel as DDP

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from common import setup  # noqa: E402


class Block(nn.Module):
    def __init__(self, d, h):
        super().__init__()
        self.ln1, self.ln2 = nn.LayerNorm(d), nn.LayerNorm(d)
        self.qkv, self.proj = nn.Linear(d, 3 * d), nn.Linear(d, d)
        self.mlp = nn.Sequential(nn.Linear(d, 4 * d), nn.GELU(), nn.Linear(4 * d, d))
        self.h = h

    def forward(self, x):
        B, S, D = x.shape
        q, k, v = self.qkv(self.ln1(x)).view(B, S, 3, self.h, D // self.h).permute(2, 0, 3, 1, 4)
        x = x + self.proj(F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, S, D))
        return x + self.mlp(self.ln2(x))


class MiniGPT(nn.Module):
    def __init__(self, vocab, d=128, layers=3, heads=4, block=64):
        super().__init__()
        self.tok, self.pos = nn.Embedding(vocab, d), nn.Embedding(block, d)
        self.blocks = nn.Sequential(*[Block(d, heads) for _ in range(layers)])
        self.ln, self.head = nn.LayerNorm(d), nn.Linear(d, vocab, bias=False)

    def forward(self, idx):
        x = self.tok(idx) + self.pos(torch.arange(idx.shape[1], device=idx.device))
        return self.head(self.ln(self.blocks(x)))


def corpus() -> str:
    words = "the quick brown fox jumps over the lazy dog while gradients flow across eight gpus in parallel".split()
    g = torch.Generator().manual_seed(0)
    return " ".join(words[i] for i in torch.randint(0, len(words), (20000,), generator=g).tolist())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--save-every", type=int, default=10)
    ap.add_argument("--snapshot", default="/tmp/minigpt_snapshot.pt")
    ap.add_argument("--batch-size", type=int, default=16)
    a = ap.parse_args()
    rank, world, dev = setup()
    text = corpus()
    chars = sorted(set(text))
    data = torch.tensor([chars.index(c) for c in text])
    torch.manual_seed(0)
    model = MiniGPT(len(chars)).to(dev)
    opt = torch.optim.AdamW(model.parameters(), lr=3e-3)
    start = 0
    if os.path.exists(a.snapshot):                                  # every rank loads the same snapshot
        snap = torch.load(a.snapshot, map_location=dev, weights_only=False)
        model.load_state_dict(snap["model"])
        opt.load_state_dict(snap["optim"])
        start = snap["step"]
        if rank == 0:
            print(f"resuming from snapshot at step {start}", flush=True)
    model = DDP(model, device_ids=[dev.index] if dev.type == "cuda" else None)
    block = 64
    for step in range(start, a.steps):
        g = torch.Generator().manual_seed(step * world + rank)      # data order is a function of the step: resume is exact
        ix = torch.randint(0, len(data) - block - 1, (a.batch_size,), generator=g)
        x = torch.stack([data[i:i + block] for i in ix]).to(dev)
        y = torch.stack([data[i + 1:i + block + 1] for i in ix]).to(dev)
        loss = F.cross_entropy(model(x).flatten(0, 1), y.flatten())
        opt.zero_grad()
        loss.backward()
        opt.step()
        if rank == 0 and (step + 1) % 10 == 0:
            print(f"step {step + 1} loss {loss.item():.4f}", flush=True)
        if (step + 1) % a.save_every == 0:
            if rank == 0:
                tmp = a.snapshot + ".tmp"
                torch.save({"model": model.module.state_dict(), "optim": opt.state_dict(), "step": step + 1}, tmp)
                os.replace(tmp, a.snapshot)                       # atomic: a crash never leaves a torn snapshot
            dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
