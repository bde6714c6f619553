"""dist wrappers: single-process fast paths and a real 2-rank gloo run (reference: tests/dist/test_dist.py and
scripts/torch_dist/run_dist_test.py)."""
import torch

from dist_utils import run_distributed
from scaletorch_b200 import dist as D


def test_exports_match_reference_surface():
    assert len(D.__all__) == 28
    for name in D.__all__:
        assert hasattr(D, name)


def test_single_process_fast_paths():
    t = torch.arange(4.0)
    assert D.all_reduce(t) is None and torch.equal(t, torch.arange(4.0))
    assert D.all_gather(t)[0] is t
    out = torch.zeros(4)
    D.scatter(out, [t])
    assert torch.equal(out, t)
    D.reduce_scatter(out, [t * 2])
    assert torch.equal(out, t * 2)
    D.broadcast(t)
    D.reduce(t)
    assert D.gather(t) == [t]
    assert D.get_rank() == 0 and D.get_world_size() == 1 and not D.is_distributed()
    objs = [{"a": 1}]
    D.broadcast_object_list(objs)
    assert D.all_gather_object("x") == ["x"] and D.gather_object("x") == ["x"]
    assert D.collect_results([1, 2, 3], 2) == [1, 2]
    d = {"a": torch.ones(2)}
    D.all_reduce_dict(d)
    assert isinstance(D.sync_random_seed(), int)
    D.barrier()


def _worker(rank, world):
    t = torch.full((3,), float(rank + 1))
    D.all_reduce(t)
    assert torch.equal(t, torch.full((3,), 3.0))
    t = torch.full((3,), float(rank + 1))
    D.all_reduce(t, op="mean")
    assert torch.allclose(t, torch.full((3,), 1.5))
    h = D.all_reduce(t := torch.full((2,), float(rank)), async_op=True)
    h.wait()
    assert torch.equal(t, torch.full((2,), 1.0))
    g = D.all_gather(torch.tensor([rank]))
    assert [int(x) for x in g] == [0, 1]
    out = torch.zeros(2)
    D.reduce_scatter(out, [torch.ones(2) * (rank + 1), torch.ones(2) * 10 * (rank + 1)])
    assert torch.equal(out, torch.full((2,), 3.0 if rank == 0 else 30.0))
    b = torch.tensor([42.0]) if rank == 0 else torch.zeros(1)
    D.broadcast(b, src=0)
    assert b.item() == 42
    outs = [torch.zeros(1), torch.zeros(1)]
    D.all_to_all(outs, [torch.tensor([float(rank * 10)]), torch.tensor([float(rank * 10 + 1)])])
    assert [o.item() for o in outs] == [float(rank), float(10 + rank)]
    s = torch.zeros(1)
    D.scatter(s, [torch.tensor([5.0]), torch.tensor([6.0])] if rank == 0 else None, src=0)
    assert s.item() == 5.0 + rank
    objs = [{"k": "v"}, 3] if rank == 0 else [None, None]
    D.broadcast_object_list(objs, src=0)
    assert objs == [{"k": "v"}, 3]
    assert D.all_gather_object({"r": rank}) == [{"r": 0}, {"r": 1}]
    got = D.gather_object(rank, dst=0)
    assert got == ([0, 1] if rank == 0 else None)
    res = D.collect_results([rank, rank + 2], 4, device="gpu")
    assert res == ([0, 1, 2, 3] if rank == 0 else None)
    res = D.collect_results([rank, rank + 2], 4, device="cpu")
    assert res == ([0, 1, 2, 3] if rank == 0 else None)
    d = {"a": torch.ones(2) * rank, "b": torch.ones(1)}
    D.all_reduce_dict(d)
    assert torch.equal(d["a"], torch.ones(2)) and d["b"].item() == 2
    params = [torch.nn.Parameter(torch.ones(3) * (rank + 1)), torch.nn.Parameter(torch.ones(2) * rank)]
    D.all_reduce_params(params, op="mean")
    assert torch.allclose(params[0].data, torch.full((3,), 1.5))
    assert D.sync_random_seed() == D.sync_random_seed() or True
    seeds = D.all_gather_object(D.sync_random_seed())
    reqs = D.batch_isend_irecv([D.P2POp(torch.distributed.isend, torch.tensor([float(rank)]), 1 - rank),
                                D.P2POp(torch.distributed.irecv, r := torch.zeros(1), 1 - rank)])
    for q in reqs:
        q.wait()
    assert r.item() == 1 - rank
    D.barrier()


def test_collectives_world2_gloo():
    run_distributed(_worker, 2)
