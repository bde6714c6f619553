"""State machines of the communication helpers, on CPU: gradient buckets (``parallel/data_parallel/bucket.py``), the
context-parallel ring communicator (``context_parallel/cp_comms.py``) and the pipeline p2p vocabulary
(``pipeline_parallel/pp_comms.py``).  Reference suites: ``tests/parallel/test_bucket.py``, ``test_cp_comms.py``,
``test_pp_comms.py``.  Single-process checks cover bookkeeping and misuse errors; three-rank gloo runs cover the wire.
"""
import pytest
import torch
import torch.nn as nn

from dist_utils import run_distributed
from scaletorch_b200.parallel.data_parallel.bucket import BucketManager
from scaletorch_b200.runtime.flat_params import FlatParamStore


class _Net(nn.Module):
    """Parameter sizes chosen so that a 1000-element bucket limit splits them unevenly; ``big`` exceeds the limit."""

    def __init__(self):
        super().__init__()
        self.a = nn.Linear(20, 30, bias=True)       # 600 + 30
        self.b = nn.Linear(30, 10, bias=False)      # 300
        self.big = nn.Linear(40, 50, bias=False)    # 2000  (> bucket size: a bucket of its own)
        self.c = nn.Linear(10, 5, bias=True)        # 50 + 5

    def forward(self, x):
        return self.c(self.b(self.a(x))).sum() + self.big.weight.sum()


def _store(seed=0):
    torch.manual_seed(seed)
    net = _Net()
    return net, FlatParamStore(net)


# --------------------------------------------------------------------------------------------------- buckets
def test_bucket_partitioning():
    net, store = _store()
    bm = BucketManager(store, dense_group=None, bucket_size=1000)
    # every parameter in exactly one bucket
    assert sorted(id(p) for b in bm.buckets for p in b.params) == sorted(id(s.param) for s in store.slots)
    assert all(bm.param_to_bucket[id(s.param)] in bm.buckets for s in store.slots)
    # at most bucket_size elements unless a single parameter is larger than that
    for b in bm.buckets:
        n = sum(s.numel for s in b.slots)
        assert n <= 1000 or len(b.slots) == 1
    assert any(len(b.slots) == 1 and b.slots[0].numel == 2000 for b in bm.buckets)
    # buckets are filled in REVERSE registration order (the order backward produces gradients): the first bucket holds
    # the last-registered parameters
    last_registered = store.slots[-1].param
    assert any(p is last_registered for p in bm.buckets[0].params)
    # ranges are disjoint views of the flat gradient buffer
    spans = sorted((b.start, b.end) for b in bm.buckets)
    for (s0, e0), (s1, _e1) in zip(spans, spans[1:]):
        assert e0 <= s1
    for b in bm.buckets:
        assert b.grad_data.data_ptr() == store.grad_flat[b.start:b.end].data_ptr() and b.numel == b.end - b.start


def test_bucket_size_one_bucket_and_many():
    _, store = _store()
    assert len(BucketManager(store, None, bucket_size=10**9).buckets) == 1
    assert len(BucketManager(store, None, bucket_size=1).buckets) == len(store.slots)


def test_bucket_ready_bookkeeping_and_misuse():
    _, store = _store()
    bm = BucketManager(store, None, bucket_size=1000)
    b = max(bm.buckets, key=lambda x: len(x.params))
    assert len(b.params) >= 2 and not b.all_ready()
    for p in b.params[:-1]:
        bm.mark_param_as_ready(p)
        assert not b.launched
    with pytest.raises(RuntimeError, match="ready twice"):
        bm.mark_param_as_ready(b.params[0])
    bm.mark_param_as_ready(b.params[-1])            # last one launches the reduction (a no-op at world size 1)
    assert b.all_ready() and b.launched and b.handle is None
    with pytest.raises(RuntimeError, match="launched twice"):
        b.sync_gradient()
    b.wait()                                        # nothing in flight: returns
    bm.reset()
    assert not b.launched and not b.all_ready()
    # sync=False only records readiness (gradient accumulation: no reduction until the last micro-batch)
    for p in b.params:
        bm.mark_param_as_ready(p, sync=False)
    assert b.all_ready() and not b.launched
    bm.sync_all()
    assert all(x.launched for x in bm.buckets)
    bm.sync_all()                                   # idempotent: launched buckets are skipped
    bm.wait()
    bm.reset()


def test_bucket_skip_sync_flag():
    _, store = _store()
    bm = BucketManager(store, None, bucket_size=1000)
    b = bm.buckets[0]
    b.skip_sync = True
    b.sync_gradient()
    assert b.launched and b.handle is None


def _w_bucket_allreduce(rank, world):
    import torch.distributed as dist

    net, store = _store(seed=0)                     # same weights everywhere
    bm = BucketManager(store, dense_group=dist.group.WORLD, bucket_size=1000)
    torch.manual_seed(100 + rank)
    x = torch.randn(4, 20)
    loss = net(x)
    grads = torch.autograd.grad(loss, [s.param for s in store.slots], allow_unused=True)
    skip = store.slots[1].param                      # pretend one parameter got no gradient: it must contribute zeros
    for s, g in zip(store.slots, grads):
        if s.param is skip:
            s.param.main_grad.fill_(123.0)           # stale content that has to be cleared
            s.param.grad_written = False
            continue
        s.param.main_grad.copy_(g.reshape(s.param.main_grad.shape))
        s.param.grad_written = True
    for s in reversed(store.slots):
        bm.mark_param_as_ready(s.param)
    assert all(b.launched for b in bm.buckets)
    bm.wait()
    # oracle: mean over ranks of the per-rank gradients
    for s, g in zip(store.slots, grads):
        mine = torch.zeros_like(g) if s.param is skip else g
        parts = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(parts, mine.contiguous())
        want = torch.stack(parts).mean(0)
        assert torch.allclose(s.param.main_grad.reshape(want.shape), want, atol=1e-6), s.name
    bm.reset()
    dist.barrier()


def test_bucket_allreduce_three_ranks():
    run_distributed(_w_bucket_allreduce, 3)


# --------------------------------------------------------------------------------------------------- CP ring
def test_cp_communicator_single_process_and_misuse():
    from scaletorch_b200.parallel.context_parallel.cp_comms import ContextCommunicator

    c = ContextCommunicator("kv")
    assert c.world_size == 1
    x = torch.arange(6.0).view(2, 3)
    y = c.send_recv(x)                               # ring of one: receive what was sent
    assert torch.equal(y, x) and y.data_ptr() != x.data_ptr()
    buf = torch.zeros(2, 3)
    assert c.send_recv(x.t().contiguous().t(), buf) is buf and torch.equal(buf, x)
    with pytest.raises(RuntimeError, match="before commit"):
        c.wait()
    c.commit()
    with pytest.raises(RuntimeError, match="twice"):
        c.commit()
    with pytest.raises(RuntimeError, match="after commit"):
        c.send_recv(x)
    c.wait()
    c.commit()                                       # reusable after wait()
    c.wait()


def _w_cp_ring(rank, world):
    import torch.distributed as dist

    from scaletorch_b200.parallel.context_parallel.cp_comms import ContextCommunicator

    c = ContextCommunicator("kv", group=dist.group.WORLD, send_rank=(rank + 1) % world, recv_rank=(rank - 1) % world)
    assert c.world_size == world
    k = torch.full((2, 4), float(rank))
    v = torch.full((3,), float(10 + rank))
    for step in range(1, world + 1):                 # a full revolution brings the own block back
        nk, nv = c.send_recv(k), c.send_recv(v)      # two tensors batched into one commit
        c.commit()
        c.wait()
        k, v = nk, nv
        src = (rank - step) % world
        assert torch.equal(k, torch.full((2, 4), float(src))) and torch.equal(v, torch.full((3,), float(10 + src)))
    # a non-contiguous send buffer is staged, and the staging copy survives until wait()
    nc = torch.arange(8.0).view(2, 4).t() + rank
    got = c.send_recv(nc)
    c.commit()
    c.wait()
    assert torch.equal(got, torch.arange(8.0).view(2, 4).t() + (rank - 1) % world)
    dist.barrier()


def test_cp_ring_three_ranks():
    run_distributed(_w_cp_ring, 3)


# --------------------------------------------------------------------------------------------------- PP p2p
def test_pp_comms_argument_errors():
    from scaletorch_b200.parallel.pipeline_parallel import pp_comms

    with pytest.raises(ValueError, match="Invalid operation"):
        pp_comms.pipeline_communicate("send_sideways")
    with pytest.raises(ValueError, match="Invalid operation"):
        pp_comms.bidirectional_pipeline_communicate("send_fwd_recv_fwd", torch.zeros(1), (1,), "cpu", torch.float32)
    with pytest.raises(ValueError, match="Invalid recv operation"):
        pp_comms.post_recv("send_forward", (1,), "cpu", torch.float32)


def _w_pp_chain(rank, world):
    import torch.distributed as dist

    from scaletorch_b200.parallel.pipeline_parallel import pp_comms
    from scaletorch_b200.parallel.process_group import setup_process_group_manager

    pgm = setup_process_group_manager(tp_size=1, cp_size=1, pp_size=world, dp_size=1)
    assert pgm.pp_rank == rank and pgm.pp_is_first_stage == (rank == 0) and pgm.pp_is_last_stage == (rank == world - 1)
    shape, kw = (2, 3), dict(device="cpu", dtype=torch.float32)
    with pytest.raises(ValueError, match="shapes must be provided"):
        pp_comms.pipeline_communicate("recv_forward", **kw)
    # forward chain: every stage adds its rank
    x = pp_comms.pipeline_communicate("recv_forward", shapes=shape, **kw)
    assert (x is None) == (rank == 0)
    if x is not None:
        assert x.requires_grad                       # received activations are autograd leaves
    h = (torch.zeros(shape) if x is None else x.detach()) + rank
    assert pp_comms.pipeline_communicate("send_forward", tensor=h, **kw) is None
    assert torch.equal(h, torch.full(shape, float(sum(range(rank + 1)))))
    # backward chain, receive pre-posted before the "compute"
    pend = pp_comms.post_recv("recv_backward", shape, "cpu", torch.float32)
    assert (pend is None) == (rank == world - 1)
    g = torch.ones(shape) if pend is None else pend.wait() * 2
    assert not g.requires_grad
    pp_comms.pipeline_communicate("send_backward", tensor=g, **kw)
    assert torch.equal(g, torch.full(shape, float(2 ** (world - 1 - rank))))
    if rank != 0:
        with pytest.raises(ValueError, match="tensor must be provided"):
            pp_comms.pipeline_communicate("send_backward", **kw)
    dist.barrier()
    # 1F1B steady state: send activation forward and receive its gradient from the same neighbour in one call
    if rank < world - 1:
        got = pp_comms.bidirectional_pipeline_communicate("send_fwd_recv_bwd", torch.full(shape, float(rank)), shape, "cpu",
                                                          torch.float32)
        assert torch.equal(got, torch.full(shape, float(100 + rank + 1))) and not got.requires_grad
    if rank > 0:
        got = pp_comms.bidirectional_pipeline_communicate("send_bwd_recv_fwd", torch.full(shape, float(100 + rank)), shape,
                                                          "cpu", torch.float32)
        assert torch.equal(got, torch.full(shape, float(rank - 1))) and got.requires_grad
    # edges: the last stage has nobody to send forward to, the first nobody to send backward to
    if rank == world - 1:
        assert pp_comms.bidirectional_pipeline_communicate("send_fwd_recv_bwd", torch.zeros(shape), shape, "cpu",
                                                           torch.float32) is None
    if rank == 0:
        assert pp_comms.bidirectional_pipeline_communicate("send_bwd_recv_fwd", torch.zeros(shape), shape, "cpu",
                                                           torch.float32) is None
    dist.barrier()


def test_pp_chain_three_stages():
    run_distributed(_w_pp_chain, 3)
