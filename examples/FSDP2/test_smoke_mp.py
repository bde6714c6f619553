"""Multi-process smoke test of the FSDP2 examples without torchrun: ``mp.spawn`` + gloo on CPU (reference:
``examples/FSDP2/test_smoke_mp.py``).  Checks that (1) FSDP2 over 2 ranks reproduces the single-process loss trajectory of
the same model on the same global batch, (2) a distributed checkpoint written mid-run resumes to the same parameters.

    python examples/FSDP2/test_smoke_mp.py        (or: pytest examples/FSDP2/test_smoke_mp.py)
"""
from __future__ import annotations

import os
import socket
import sys
import tempfile

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def _port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _batch(step: int, vocab: int, batch: int, seq: int):
    g = torch.Generator().manual_seed(77 + step)
    tok = torch.randint(0, vocab, (batch, seq + 1), generator=g)
    return tok[:, :-1], tok[:, 1:]


def _train(model, opt, steps, rank, world, vocab, start=0, global_batch=8):
    losses = []
    for step in range(start, steps):
        x, y = _batch(step, vocab, global_batch, 32)
        x, y = x[rank::world], y[rank::world]                  # disjoint shares of one global batch
        loss = F.cross_entropy(model(x).float().flatten(0, 1), y.flatten())
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        t = loss.detach().clone()
        if world > 1:
            dist.all_reduce(t)
            t /= world
        losses.append(t.item())
    return losses


def _worker(rank: int, world: int, port: int, tmp: str, ref_losses):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from torch.distributed.fsdp import fully_shard

        from checkpoint import load_checkpoint, save_checkpoint
        from llama2 import Llama2, Llama2Config

        cfg = Llama2Config.named("tiny")

        def fresh():
            torch.manual_seed(0)
            m = Llama2(cfg)
            m.init_weights()
            for blk in m.layers:
                fully_shard(blk)
            fully_shard(m)
            return m, torch.optim.SGD(m.parameters(), lr=0.1)

        model, opt = fresh()
        losses = _train(model, opt, 4, rank, world, cfg.vocab_size)
        for a, b in zip(losses, ref_losses):
            assert abs(a - b) < 2e-4 * max(1.0, abs(b)), (losses, ref_losses)
        save_checkpoint(model, opt, tmp, 4)
        _train(model, opt, 6, rank, world, cfg.vocab_size, start=4)
        want = [p.full_tensor().clone() for p in model.parameters()]
        model2, opt2 = fresh()
        assert load_checkpoint(model2, opt2, tmp) == 4
        _train(model2, opt2, 6, rank, world, cfg.vocab_size, start=4)
        for p, w in zip(model2.parameters(), want):
            assert torch.allclose(p.full_tensor(), w, atol=1e-6), "resumed run diverged from the uninterrupted one"
        if rank == 0:
            print("FSDP2 mp smoke test: ok", flush=True)
    finally:
        dist.destroy_process_group()


def test_fsdp2_matches_single_process_and_resumes():
    from llama2 import Llama2, Llama2Config

    cfg = Llama2Config.named("tiny")
    torch.manual_seed(0)
    ref = Llama2(cfg)
    ref.init_weights()
    ref_losses = _train(ref, torch.optim.SGD(ref.parameters(), lr=0.1), 4, 0, 1, cfg.vocab_size)
    # the single process sees the whole global batch of the 2-rank run (8 rows): same mean loss, same gradient
    with tempfile.TemporaryDirectory() as tmp:
        mp.spawn(_worker, args=(2, _port(), tmp, ref_losses), nprocs=2, join=True)


if __name__ == "__main__":
    test_fsdp2_matches_single_process_and_resumes()
