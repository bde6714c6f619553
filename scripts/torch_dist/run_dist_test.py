#!/usr/bin/env python
"""Self-checks of the ``scaletorch_b200.dist`` API and of the peer-memory kernels on a *real* back-end.

Run under torchrun:   torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 8 scripts/torch_dist/run_dist_test.py
(CPU: add ``--backend gloo``).  Counterpart of the reference's ``scripts/torch_dist/run_dist_test.py:14-554`` (14
collective checks under NCCL/HCCL/gloo); on NCCL this version additionally checks the on-device barrier, the two-shot
all-reduce, the GEMM→reduce-scatter and all-gather→GEMM kernels against the c10d results.
"""
from __future__ import annotations

import argparse
import os
import sys
import traceback

import torch
import torch.distributed as tdist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from scaletorch_b200 import dist  # noqa: E402

CHECKS = []


def check(fn):
    CHECKS.append(fn)
    return fn


def dev():
    return torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0))) if tdist.get_backend() == "nccl" else torch.device("cpu")


@check
def all_reduce_sum_mean_max():
    r, w = dist.get_rank(), dist.get_world_size()
    for op, want in (("sum", w * (w - 1) / 2), ("mean", (w - 1) / 2), ("max", w - 1), ("min", 0)):
        t = torch.full((1024,), float(r), device=dev())
        dist.all_reduce(t, op)
        assert torch.allclose(t, torch.full_like(t, want)), (op, t[0].item(), want)


@check
def all_reduce_async():
    t = torch.ones(4096, device=dev())
    h = dist.all_reduce(t, "sum", async_op=True)
    if h is not None:
        h.wait()
    assert t[0].item() == dist.get_world_size()


@check
def all_gather():
    r, w = dist.get_rank(), dist.get_world_size()
    out = dist.all_gather(torch.full((3, 5), float(r), device=dev()))
    assert len(out) == w and all(out[i][0, 0].item() == i for i in range(w))


@check
def gather():
    r, w = dist.get_rank(), dist.get_world_size()
    out = dist.gather(torch.tensor([float(r)], device=dev()), dst=0)
    if r == 0:
        assert [o.item() for o in out] == list(range(w))
    else:
        assert out == []


@check
def broadcast():
    t = torch.arange(16.0, device=dev()) if dist.get_rank() == 0 else torch.zeros(16, device=dev())
    dist.broadcast(t, src=0)
    assert torch.equal(t.cpu(), torch.arange(16.0))


@check
def scatter():
    r, w = dist.get_rank(), dist.get_world_size()
    out = torch.empty(4, device=dev())
    lst = [torch.full((4,), float(i), device=dev()) for i in range(w)] if r == 0 else None
    dist.scatter(out, lst, src=0)
    assert out[0].item() == r


@check
def reduce():
    r, w = dist.get_rank(), dist.get_world_size()
    t = torch.full((8,), float(r + 1), device=dev())
    dist.reduce(t, dst=0, op="sum")
    if r == 0:
        assert t[0].item() == w * (w + 1) / 2


@check
def reduce_scatter():
    r, w = dist.get_rank(), dist.get_world_size()
    out = torch.empty(6, device=dev())
    dist.reduce_scatter(out, [torch.full((6,), float(i + r), device=dev()) for i in range(w)])
    assert out[0].item() == w * r + w * (w - 1) / 2, out


@check
def all_to_all():
    r, w = dist.get_rank(), dist.get_world_size()
    ins = [torch.full((2,), float(r * 100 + i), device=dev()) for i in range(w)]
    outs = [torch.empty(2, device=dev()) for _ in range(w)]
    dist.all_to_all(outs, ins)
    assert [o[0].item() for o in outs] == [float(i * 100 + r) for i in range(w)]


@check
def batch_isend_irecv_ring():
    r, w = dist.get_rank(), dist.get_world_size()
    if w < 2:
        return
    send = torch.full((32,), float(r), device=dev())
    recv = torch.empty(32, device=dev())
    ops = [dist.P2POp(tdist.isend, send, (r + 1) % w), dist.P2POp(tdist.irecv, recv, (r - 1) % w)]
    for h in dist.batch_isend_irecv(ops):
        h.wait()
    if recv.is_cuda:
        torch.cuda.synchronize()
    assert recv[0].item() == (r - 1) % w


@check
def object_collectives():
    r, w = dist.get_rank(), dist.get_world_size()
    objs = dist.all_gather_object({"rank": r, "name": f"r{r}"})
    assert [o["rank"] for o in objs] == list(range(w))
    lst = [{"cfg": 42}, "x"] if r == 0 else [None, None]
    dist.broadcast_object_list(lst, src=0)
    assert lst[0] == {"cfg": 42} and lst[1] == "x"
    g = dist.gather_object(r * 2, dst=0)
    if r == 0:
        assert g == [2 * i for i in range(w)]


@check
def all_reduce_dict_and_params():
    w = dist.get_world_size()
    d = {"a": torch.ones(3, device=dev()), "b": torch.full((2,), 2.0, device=dev())}
    dist.all_reduce_dict(d, "sum")
    assert d["a"][0].item() == w and d["b"][0].item() == 2 * w
    params = [torch.ones(5, device=dev()), torch.ones(7, device=dev())]
    dist.all_reduce_params(params, coalesce=True, op="mean")
    assert all(torch.allclose(p, torch.ones_like(p)) for p in params)


@check
def sync_seed_and_collect():
    seeds = dist.all_gather_object(dist.sync_random_seed())
    assert len(set(seeds)) == 1
    r, w = dist.get_rank(), dist.get_world_size()
    part = [r + w * i for i in range(3)]
    res = dist.collect_results(part, 3 * w, device="cpu")
    if r == 0:
        assert res == list(range(3 * w)), res


@check
def sub_groups():
    r, w = dist.get_rank(), dist.get_world_size()
    if w < 4 or w % 2:
        return
    groups = [dist.new_group(list(range(i, i + 2))) for i in range(0, w, 2)]
    g = groups[r // 2]
    t = torch.full((4,), float(r), device=dev())
    dist.all_reduce(t, "sum", group=g)
    assert t[0].item() == (r // 2) * 4 + 1
    assert dist.get_world_size(g) == 2 and dist.get_rank(g) == r % 2


@check
def peer_memory_kernels():
    """NCCL only: on-device barrier / all-reduce / GEMM→RS / AG→GEMM over symmetric memory vs c10d."""
    if tdist.get_backend() != "nccl":
        return
    from scaletorch_b200.runtime import fused_ops as F
    from scaletorch_b200.runtime import symm

    grp = tdist.group.WORLD
    if not symm.group_ready(grp):
        raise AssertionError("symmetric memory could not be set up on this fabric")
    r, w = dist.get_rank(), dist.get_world_size()
    g = F.get_group(grp)
    g.barrier()
    t = torch.randn(1 << 20, device=dev()).bfloat16()
    ref = t.clone()
    tdist.all_reduce(ref)
    assert (F.all_reduce_(t, grp).float() - ref.float()).norm() / ref.float().norm() < 1e-2
    x = (torch.randn(256 * w, 1024, device=dev()) * 0.5).bfloat16()
    wt = (torch.randn(768, 1024, device=dev()) * 0.05).bfloat16()
    full = x.float() @ wt.float().t()
    tdist.all_reduce(full)
    y = F.gemm_reduce_scatter(x, wt, grp)
    want = full[r * 256:(r + 1) * 256]
    assert (y.float() - want).norm() / want.norm() < 2e-2
    xs = (torch.randn(256, 1024, device=dev()) * 0.5).bfloat16()
    gathered = torch.empty(256 * w, 1024, device=dev(), dtype=torch.bfloat16)
    tdist.all_gather_into_tensor(gathered, xs)
    y2, xf = F.all_gather_gemm(xs, wt, grp)
    assert torch.equal(xf, gathered)
    want2 = gathered.float() @ wt.float().t()
    assert (y2.float() - want2).norm() / want2.norm() < 2e-2


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default=None)
    a = ap.parse_args()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_dist("pytorch", backend=a.backend)
    rank = dist.get_rank()
    failed = 0
    for fn in CHECKS:
        try:
            fn()
            ok = 1
        except Exception:
            ok = 0
            print(f"[rank {rank}] {fn.__name__} FAILED\n{traceback.format_exc()}", flush=True)
        flags = dist.all_gather_object(ok)
        if rank == 0:
            print(f"{'PASS' if all(flags) else 'FAIL'}  {fn.__name__}", flush=True)
        failed += 0 if all(flags) else 1
    if rank == 0:
        print(f"{len(CHECKS) - failed}/{len(CHECKS)} checks passed on backend {tdist.get_backend()} with world size {dist.get_world_size()}")
    dist.cleanup_dist()
    return 1 if failed else 0


if __name__ == "__main__":
    sys.exit(main())
