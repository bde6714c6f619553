"""One test per wrapper of ``scaletorch_b200.dist`` (the reference's 28 exported helpers,
``scaletorch/dist/__init__.py``; its own suite is ``tests/dist/test_dist.py``): helpers and single-process behaviour first,
then a THREE-rank gloo run (odd world size, non-zero roots, sub-groups) so that src / dst / group-rank translation and the
gloo emulations (reduce-scatter through all-reduce, all-to-all through p2p) are exercised where a 2-rank run hides bugs.
"""
import pytest
import torch

from dist_utils import run_distributed
from scaletorch_b200 import dist as D
from scaletorch_b200.dist import collective_ops as C
from scaletorch_b200.dist import gather_utils as G
from scaletorch_b200.dist import object_ops as O
from scaletorch_b200.dist import p2p_ops as P
from scaletorch_b200.dist import utils as U
from scaletorch_b200.dist._reduce_op import _get_reduce_op


# ------------------------------------------------------------------------------------------ helpers, no process group
def test_reduce_op_names():
    import torch.distributed as td

    assert _get_reduce_op("sum") == td.ReduceOp.SUM and _get_reduce_op("MEAN") == td.ReduceOp.SUM
    assert _get_reduce_op("max") == td.ReduceOp.MAX and _get_reduce_op("min") == td.ReduceOp.MIN
    assert _get_reduce_op("product") == td.ReduceOp.PRODUCT
    for name in ("band", "bor", "bxor"):
        _get_reduce_op(name)
    with pytest.raises(ValueError, match="reduce op should be one of"):
        _get_reduce_op("avg")


def test_flatten_unflatten_roundtrip():
    ts = [torch.arange(6.0).view(2, 3), torch.tensor(7.0), torch.arange(4.0).view(2, 2, 1)]
    flat = C._flatten_dense_tensors(ts)
    assert flat.shape == (11,)
    back = C._unflatten_dense_tensors(flat, ts)
    for a, b in zip(ts, back):
        assert a.shape == b.shape and torch.equal(a, b)
    back[0].add_(1)                                            # views of the flat buffer, not copies
    assert flat[0].item() == 1.0
    assert C._flatten_dense_tensors([]).numel() == 0


def test_take_tensors_respects_limit_and_dtype():
    a = [torch.zeros(100), torch.zeros(100), torch.zeros(100)]          # 400 B each
    i = [torch.zeros(10, dtype=torch.int64)]
    buckets = list(C._take_tensors(a + i, 900))
    sizes = sorted(sum(t.numel() * t.element_size() for t in b) for b in buckets)
    assert sizes == [80, 400, 800]
    for b in buckets:
        assert len({t.dtype for t in b}) == 1
    # a single tensor larger than the limit still travels (alone)
    big = list(C._take_tensors([torch.zeros(1000)], 16))
    assert len(big) == 1 and big[0][0].numel() == 1000


def test_get_data_device_and_cast():
    t = torch.zeros(2)
    assert U.get_data_device(t) == torch.device("cpu")
    assert U.get_data_device([t, t]) == torch.device("cpu")
    assert U.get_data_device({"a": t, "b": [t]}) == torch.device("cpu")
    with pytest.raises((TypeError, ValueError)):
        U.get_data_device("not a tensor")
    same = U.cast_data_device(t, torch.device("cpu"))
    assert same is t                                            # no copy when already there
    out = torch.ones(2)
    U.cast_data_device(torch.full((2,), 3.0), torch.device("cpu"), out=out)
    assert torch.equal(out, torch.full((2,), 3.0))
    nested = U.cast_data_device({"a": t, "b": [t, t]}, torch.device("cpu"))
    assert set(nested) == {"a", "b"} and len(nested["b"]) == 2


def test_infer_launcher_from_env(monkeypatch):
    for k in ("WORLD_SIZE", "SLURM_NTASKS", "OMPI_COMM_WORLD_LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    assert U.infer_launcher() == "none"
    monkeypatch.setenv("OMPI_COMM_WORLD_LOCAL_RANK", "0")
    assert U.infer_launcher() == "mpi"
    monkeypatch.setenv("SLURM_NTASKS", "4")
    assert U.infer_launcher() == "slurm"
    monkeypatch.setenv("WORLD_SIZE", "2")
    assert U.infer_launcher() == "pytorch"


def test_init_dist_rejects_unknown_launcher():
    with pytest.raises(ValueError):
        U.init_dist("carrier-pigeon")


def test_rank_helpers_without_process_group():
    assert not U.is_distributed()
    assert U.get_rank() == 0 and U.get_world_size() == 1 and U.get_dist_info() == (0, 1)
    assert U.get_local_rank() == 0 and U.get_local_size() == 1
    assert U.is_main_process() and U.get_backend() is None
    assert U.get_default_group() is None and U.get_local_group() is None
    assert U.get_comm_device() == torch.device("cpu")
    U.barrier()
    U.cleanup_dist()                                            # idempotent without a group

    calls = []

    @U.master_only
    def f(x):
        calls.append(x)
        return x + 1

    assert f(1) == 2 and calls == [1] and f.__name__ == "f"


def test_interleave_and_collect_single():
    # rank r holds samples r, r + world, ...: the merge restores dataset order and drops the padding
    assert G._interleave([[0, 3, 6], [1, 4, 7], [2, 5, 8]], 8) == list(range(8))
    assert G.collect_results([1, 2, 3], 3) == [1, 2, 3]
    assert G.collect_results([1, 2, 3], 2, device="gpu") == [1, 2]
    with pytest.raises((ValueError, NotImplementedError)):
        G.collect_results([1], 1, device="tpu")


def test_object_tensor_roundtrip():
    obj = {"a": [1, 2, 3], "b": "text", "c": torch.arange(3)}
    t, n = O._object_to_tensor(obj)
    assert t.dtype == torch.uint8 and int(n.item()) == t.numel()
    back = O._tensor_to_object(t, int(n.item()))
    assert back["a"] == [1, 2, 3] and back["b"] == "text" and torch.equal(back["c"], torch.arange(3))


def test_single_process_identities():
    t = torch.arange(5.0)
    for op in ("sum", "mean", "max", "min", "product"):
        assert D.all_reduce(t.clone(), op=op) is None
    assert D.all_reduce(t, async_op=True) is None
    assert D.all_gather(t)[0] is t and D.gather(t, dst=0) == [t]
    D.broadcast(t, src=0)
    D.reduce(t, dst=0, op="mean")
    out = torch.zeros(5)
    D.scatter(out, [t], src=0)
    assert torch.equal(out, t)
    D.scatter(out)                                              # nothing to scatter: untouched
    assert torch.equal(out, t)
    outs = [torch.zeros(2)]
    D.all_to_all(outs, [torch.ones(2)])
    assert torch.equal(outs[0], torch.ones(2))
    with pytest.raises(TypeError):
        D.all_reduce_dict([t])
    D.all_reduce_dict({})
    D.all_reduce_params(iter([torch.nn.Parameter(t.clone())]))
    objs = [1, "a"]
    D.broadcast_object_list(objs)
    assert objs == [1, "a"]
    assert D.all_gather_object(3) == [3] and D.gather_object(3) == [3]
    assert D.batch_isend_irecv([]) == []


# ------------------------------------------------------------------------------------------ three ranks, gloo
def _w_reduce_family(rank, world):
    base = torch.tensor([1.0, 2.0, 3.0]) * (rank + 1)
    for op, want in (("sum", 6.0), ("mean", 2.0), ("max", 3.0), ("min", 1.0), ("product", 6.0)):
        t = base.clone()
        D.all_reduce(t, op=op)
        assert torch.allclose(t, torch.tensor([1.0, 2.0, 3.0]) * want if op != "product" else torch.tensor([6.0, 48.0, 162.0])), op
    i = torch.tensor([rank + 1])                                # integer mean must not truncate in place
    D.all_reduce(i, op="sum")
    assert i.item() == 6
    h = D.all_reduce(t := base.clone(), op="mean", async_op=True)
    h.wait()
    assert torch.allclose(t, torch.tensor([2.0, 4.0, 6.0]))
    with pytest.raises(ValueError):
        D.all_reduce(base.clone(), op="median")
    # reduce to a NON-zero root: only the root's value is specified
    r = base.clone()
    D.reduce(r, dst=2, op="sum")
    if rank == 2:
        assert torch.allclose(r, torch.tensor([6.0, 12.0, 18.0]))
    r = base.clone()
    D.reduce(r, dst=1, op="mean")
    if rank == 1:
        assert torch.allclose(r, torch.tensor([2.0, 4.0, 6.0]))
    out = torch.zeros(2)
    D.reduce_scatter(out, [torch.full((2,), float(10 * j + rank)) for j in range(world)])
    assert torch.equal(out, torch.full((2,), float(30 * rank + 3)))
    out = torch.zeros(2)
    D.reduce_scatter(out, [torch.full((2,), float(j + rank)) for j in range(world)], op="mean")
    assert torch.allclose(out, torch.full((2,), float(rank + 1)))
    with pytest.raises(ValueError):
        D.reduce_scatter(out, [torch.zeros(2)])
    D.barrier()


def _w_gather_family(rank, world):
    x = torch.arange(4).view(2, 2) + 10 * rank
    got = D.all_gather(x)
    assert len(got) == world and all(torch.equal(g, torch.arange(4).view(2, 2) + 10 * j) for j, g in enumerate(got))
    got = D.gather(x, dst=1)
    if rank == 1:
        assert [int(g[0, 0]) for g in got] == [0, 10, 20]
    else:
        assert got == []
    b = torch.full((3,), 7.0) if rank == 2 else torch.zeros(3)
    D.broadcast(b, src=2)
    assert torch.equal(b, torch.full((3,), 7.0))
    nc = torch.arange(6.0).view(2, 3).t() * (1 if rank == 0 else 0)      # non-contiguous destination
    D.broadcast(nc, src=0)
    assert torch.equal(nc, torch.arange(6.0).view(2, 3).t())
    s = torch.zeros(2)
    D.scatter(s, [torch.full((2,), float(j)) for j in range(world)] if rank == 1 else None, src=1)
    assert torch.equal(s, torch.full((2,), float(rank)))
    if rank == 0:
        with pytest.raises(ValueError):
            D.scatter(torch.zeros(1), [torch.zeros(1)], src=0)
    D.barrier()
    outs = [torch.zeros(2) for _ in range(world)]
    D.all_to_all(outs, [torch.full((2,), float(100 * rank + j)) for j in range(world)])
    assert [int(o[0]) for o in outs] == [100 * j + rank for j in range(world)]
    with pytest.raises(ValueError):
        D.all_to_all(outs[:1], outs[:1])
    D.barrier()


def _w_objects_and_params(rank, world):
    big = {"rank": rank, "blob": list(range(1000 * (rank + 1)))}           # different pickled sizes per rank
    got = D.all_gather_object(big)
    assert [g["rank"] for g in got] == [0, 1, 2] and [len(g["blob"]) for g in got] == [1000, 2000, 3000]
    got = D.gather_object(big, dst=2)
    assert (got is None) == (rank != 2)
    if rank == 2:
        assert [len(g["blob"]) for g in got] == [1000, 2000, 3000]
    objs = [{"cfg": [1, 2]}, "s", 3.5] if rank == 1 else [None, None, None]
    D.broadcast_object_list(objs, src=1)
    assert objs == [{"cfg": [1, 2]}, "s", 3.5]
    seeds = D.all_gather_object(D.sync_random_seed())
    assert len(set(seeds)) == 1
    # results of a sharded evaluation: rank r holds samples r, r + world, ...; 7 samples, the last shard padded
    part = [rank + world * j for j in range(3)]
    for dev in ("cpu", "gpu"):
        res = D.collect_results(part, 7, device=dev)
        assert res == (list(range(7)) if rank == 0 else None)
    d = {"loss": torch.tensor([float(rank)]), "acc": torch.tensor([1.0, 2.0]) * rank}
    D.all_reduce_dict(d, op="mean")
    assert torch.allclose(d["loss"], torch.tensor([1.0])) and torch.allclose(d["acc"], torch.tensor([1.0, 2.0]))
    ps = [torch.nn.Parameter(torch.full((5,), float(rank))), torch.nn.Parameter(torch.full((3,), 2.0 * rank)),
          torch.nn.Parameter(torch.full((2,), rank, dtype=torch.float64))]
    D.all_reduce_params(ps, coalesce=True, bucket_size_mb=-1)
    assert ps[0].data[0] == 3 and ps[1].data[0] == 6 and ps[2].data[0] == 3 and ps[2].dtype == torch.float64
    ps = [torch.nn.Parameter(torch.full((300_000,), float(rank))) for _ in range(3)]      # 1.2 MB each, 1 MB buckets
    D.all_reduce_params(ps, coalesce=True, bucket_size_mb=1, op="mean")
    assert all(torch.allclose(p.data, torch.ones_like(p.data)) for p in ps)
    ps = [torch.nn.Parameter(torch.full((4,), float(rank)))]
    D.all_reduce_params((p for p in ps), coalesce=False)
    assert ps[0].data[0] == 3
    D.barrier()


def _w_groups_and_p2p(rank, world):
    assert D.is_distributed() and D.get_world_size() == world and D.get_rank() == rank
    assert U.get_dist_info() == (rank, world) and U.get_backend() == "gloo"
    assert U.is_main_process() == (rank == 0)
    assert U.get_comm_device() == torch.device("cpu")
    sub = D.new_group([0, 2])                                   # every rank calls new_group; rank 1 is not a member
    if rank in (0, 2):
        assert D.get_world_size(sub) == 2 and D.get_rank(sub) == (0 if rank == 0 else 1)
        t = torch.tensor([float(rank)])
        D.all_reduce(t, group=sub)
        assert t.item() == 2.0
        b = torch.tensor([9.0]) if rank == 2 else torch.zeros(1)
        D.broadcast(b, src=1, group=sub)                        # src is a GROUP rank: group rank 1 = global rank 2
        assert b.item() == 9.0
        got = D.all_gather(torch.tensor([rank]), group=sub)
        assert [int(g) for g in got] == [0, 2]
    D.barrier()
    # ring exchange with batched p2p: send to the right neighbour, receive from the left
    right, left = (rank + 1) % world, (rank - 1) % world
    recv = torch.zeros(2)
    ops = [D.P2POp(torch.distributed.isend, torch.full((2,), float(rank)), right),
           D.P2POp(torch.distributed.irecv, recv, left)]
    for q in D.batch_isend_irecv(ops):
        q.wait()
    assert torch.equal(recv, torch.full((2,), float(left)))
    if rank == 0:
        P.isend(torch.tensor([5.0]), 1, tag=3).wait()
    elif rank == 1:
        r = torch.zeros(1)
        P.irecv(r, 0, tag=3).wait()
        assert r.item() == 5.0
    ran = []

    @U.master_only
    def only_zero():
        ran.append(rank)

    only_zero()
    assert ran == ([0] if rank == 0 else [])
    U.init_local_group(0, world)                                # one "node" holding every rank
    assert U.get_local_size() == world and U.get_local_rank() == rank and U.get_local_group() is not None
    D.barrier()
    D.destroy_group(sub)


@pytest.mark.parametrize("worker", [_w_reduce_family, _w_gather_family, _w_objects_and_params, _w_groups_and_p2p],
                         ids=["reduce", "gather", "objects", "groups_p2p"])
def test_three_ranks_gloo(worker):
    run_distributed(worker, 3)
