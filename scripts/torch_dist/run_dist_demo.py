#!/usr/bin/env python
"""Guided tour of the communication layer, printing what every rank sees (reference: ``scripts/torch_dist/run_dist_demo.py``).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 4 scripts/torch_dist/run_dist_demo.py [--backend gloo]

Sections: (1) environment + device, (2) collectives of ``scaletorch_b200.dist``, (3) the 5-D process grid of
``ProcessGroupManager`` for a chosen layout, (4) on NCCL: bandwidth of the peer-memory all-reduce vs NCCL.
"""
from __future__ import annotations

import argparse
import os
import sys
import time

import torch
import torch.distributed as tdist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from scaletorch_b200 import dist  # noqa: E402


def rank_print(msg: str) -> None:
    if dist.get_rank() == 0:
        print(msg, flush=True)


def say(msg: str) -> None:
    """Every rank prints, one at a time, in rank order."""
    for r in range(dist.get_world_size()):
        if r == dist.get_rank():
            print(f"[rank {r}] {msg}", flush=True)
        dist.barrier()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default=None)
    ap.add_argument("--tp", type=int, default=2)
    ap.add_argument("--cp", type=int, default=1)
    ap.add_argument("--pp", type=int, default=1)
    ap.add_argument("--ep", type=int, default=1)
    a = ap.parse_args()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_dist("pytorch", backend=a.backend)
    rank, world = dist.get_rank(), dist.get_world_size()
    nccl = tdist.get_backend() == "nccl"
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))) if nccl else torch.device("cpu")

    rank_print("== 1. environment ==")
    say(f"backend={tdist.get_backend()} device={device} host={os.uname().nodename}")

    rank_print("== 2. collectives ==")
    t = torch.full((4,), float(rank), device=device)
    dist.all_reduce(t, "sum")
    say(f"all_reduce(sum) of rank ids -> {t.tolist()}")
    say(f"all_gather -> {[int(x[0]) for x in dist.all_gather(torch.tensor([rank], device=device))]}")
    outs = [torch.empty(1, device=device) for _ in range(world)]
    dist.all_to_all(outs, [torch.tensor([float(rank * 10 + i)], device=device) for i in range(world)])
    say(f"all_to_all -> {[int(o.item()) for o in outs]}")
    say(f"all_gather_object -> {dist.all_gather_object({'rank': rank})}")

    rank_print("== 3. process grid ==")
    mp = a.tp * a.cp * a.pp * a.ep
    if world % mp == 0:
        from scaletorch_b200.parallel.process_group import process_group_manager as pgm
        from scaletorch_b200.parallel.process_group import setup_process_group_manager

        setup_process_group_manager(tp_size=a.tp, cp_size=a.cp, pp_size=a.pp, dp_size=world // mp, ep_size=a.ep)
        say(str(pgm))
    else:
        rank_print(f"world size {world} is not divisible by tp*cp*pp*ep={mp}; skipping")

    if nccl:
        rank_print("== 4. peer-memory all-reduce vs NCCL (bf16, 64 MiB) ==")
        from scaletorch_b200.runtime import fused_ops as F
        from scaletorch_b200.runtime import symm

        n = 32 << 20
        x = torch.randn(n, device=device).bfloat16()
        if symm.group_ready(tdist.group.WORLD) and n * 2 <= symm.region_bytes(tdist.group.WORLD):
            for name, fn in (("nccl", lambda: tdist.all_reduce(x)), ("peer-memory", lambda: F.all_reduce_(x, tdist.group.WORLD))):
                for _ in range(3):
                    fn()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(10):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / 10
                rank_print(f"{name:12s} {ms * 1e3:8.1f} us   bus bandwidth {2 * (world - 1) / world * n * 2 / ms / 1e6:7.1f} GB/s")
        else:
            rank_print("symmetric memory unavailable on this fabric")
    time.sleep(0.1)
    dist.cleanup_dist()


if __name__ == "__main__":
    main()
