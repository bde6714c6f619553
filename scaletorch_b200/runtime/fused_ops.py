"""Python side of the fused compute+collective kernels (csrc/comm.cu, csrc/gemm_tcgen05.cu).

One :class:`SymmGroup` per process group owns
* a *flag page* per rank (monotonic epochs, 8 slots) used by the on-device cross-GPU barriers, and
* a *data window* per rank, carved into fixed regions (all-gather shards / reduce-scatter destinations /
  all-reduce partials / gradient buckets),
both allocated with ``torch.distributed._symmetric_memory`` so every peer's copy is mapped into this process
(and, when the fabric supports it, one NVLS multicast address covers all copies).  The kernels receive the
device array of peer base pointers; nothing goes through NCCL.

Ops (all stream-ordered on the current stream, no host synchronisation):
    all_gather_gemm      y = all_gather(x_shard) · Wᵀ in one kernel (copy CTAs + GEMM CTAs)      [column-parallel fwd, row-parallel bwd]
    gemm_reduce_scatter  y_shard = reduce_scatter(x · Wᵀ): epilogue TMA reduce-adds into peer memory [row-parallel fwd, column-parallel bwd]
    gemm_all_reduce      y = all_reduce(x · Wᵀ): GEMM into the window + one-shot peer reduction       [row-parallel fwd without SP]
    all_reduce_f32_      in-place mean/sum of a range of the gradient window (two-shot / NVLS)         [data-parallel buckets]
"""
from __future__ import annotations

import ctypes
import os

import torch
import torch.distributed as dist

from .. import _native as N

_groups: dict[int, "SymmGroup"] = {}
_P, _I, _L, _U, _F = ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_uint, ctypes.c_float
N.register("stb_symm_barrier", [_P, _I, _I, _I, _U, _P])
N.register("stb_allreduce_f32", [_P, _P, _P, _L, _L, _F, _I, _I, _I, _U, _I, _P])
N.register("stb_allreduce_bf16", [_P, _P, _P, _L, _L, _I, _I, _I, _U, _I, _P])
N.register("stb_allreduce_oneshot_bf16", [_P, _P, _L, _P, _L, _I, _I, _I, _U, _I, _P])
N.register("stb_allgather_pull", [_P, _P, _L, _P, _L, _L, _I, _P, _I, _I, _I, _U, _I, _P])
N.register("stb_embedding_bwd_peers", [_P, _P, _L, _L, _P, _L, _I, _I, _I, _F, _I, _I, _I, _U, _I, _P])
N.register("stb_reduce_scatter_f32", [_P, _P, _P, _L, _L, _L, _L, _F, _I, _I, _I, _U, _I, _P])
N.register("stb_adamw_shard_bcast", [_P, _P, _P, _L, _P, _P, _P, _L, _F, _F, _F, _F, _P, _F, _P, _I, _I, _I, _U, _I, _P])
N.register("stb_param_bcast", [_P, _P, _P, _L, _L, _I, _I, _I, _I, _U, _I, _P])
N.register("stb_allreduce_tiny_f32", [_P, _P, _L, _I, _I, _I, _I, _U, _P])
N.register("stb_gemm_rs_bf16", [_P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _P])
N.register("stb_ag_gemm_bf16", [_P, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _L, _P, _U, _I, _I, _I, _I, _I, _P])

# flag slots (each op family has its own so families on different streams never share an epoch counter)
SLOT_BARRIER, SLOT_AG, SLOT_RS, SLOT_AR, SLOT_DP = 0, 1, 2, 3, 4
SLOT_EMB = 7     # (5 = context-parallel ring, 6 = expert-parallel dispatch/combine)

MB = 1 << 20


def _poison() -> bool:
    import os

    return os.environ.get("SCALETORCH_B200_POISON", "0") == "1"


def _empty(shape, dtype, device) -> torch.Tensor:
    """Output buffers of the fused collectives.  Debug mode ``SCALETORCH_B200_POISON=1`` fills them with NaN first, so a
    row block a kernel failed to write (a missed flag, a wrong peer offset) shows up as NaN in the very next op instead
    of as plausible stale data (the reference has no race / ordering diagnostics)."""
    if _poison():
        return torch.full(shape, float("nan"), dtype=dtype, device=device)
    return torch.empty(shape, dtype=dtype, device=device)


def _no_multimem() -> bool:
    import os

    return os.environ.get("SCALETORCH_B200_NO_MULTIMEM", "0") == "1"


def _multicast_ptr(handle) -> int:
    """NVLS multicast address of a rendezvoused symmetric buffer (0 when the fabric has no multicast support).
    ``handle.multicast_ptr`` is the source of truth; ``has_multicast_support`` is a static method taking
    ``(device_type, device_index)`` -- calling it without arguments raises, which an earlier version of this file
    swallowed and so never used NVLS."""
    try:
        return int(getattr(handle, "multicast_ptr", 0) or 0)
    except Exception:  # pragma: no cover - very old torch
        return 0


class SymmGroup:
    """Symmetric flag pages + data window of one process group."""

    def __init__(self, group, window_bytes: int = 512 * MB):
        import torch.distributed._symmetric_memory as sm

        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        dev = torch.device("cuda", torch.cuda.current_device())
        self.device = dev
        self.flags = sm.empty(4096, dtype=torch.uint8, device=dev)
        self.flags.zero_()
        # device-driven epochs (csrc/common/symm.cuh): one counter word per slot at [144 + slot], starting at 1.  The
        # kernels read and advance it themselves, so every launch passes epoch 0 and is replayable from a CUDA graph.
        self.flags.view(torch.int32)[144:152] = 1
        self.flags_h = sm.rendezvous(self.flags, group=group)
        self.window = sm.empty(window_bytes, dtype=torch.uint8, device=dev)
        self.window_h = sm.rendezvous(self.window, group=group)
        self.window_bytes = window_bytes
        self.flag_pages_dev = int(self.flags_h.buffer_ptrs_dev)
        self.bufs_dev = int(self.window_h.buffer_ptrs_dev)
        self.peer_ptrs = [int(p) for p in self.window_h.buffer_ptrs]
        self.mc_ptr = _multicast_ptr(self.window_h)
        self.epoch = [1] * 8
        # region layout (bytes): [AG shards | RS destinations | all-reduce partials | spare]
        q = window_bytes // 4
        self.off_ag, self.off_rs, self.off_ar, self.off_spare = 0, q, 2 * q, 3 * q
        self.region_bytes = q
        self.ready_flags = torch.zeros(1 << 16, dtype=torch.int32, device=dev)   # AG->GEMM row-block flags (local)
        # GEMM->reduce-scatter destinations are double-buffered halves of the RS region; both start out zeroed and each
        # call re-zeroes the half the NEXT call will add into (see gemm_reduce_scatter)
        self.window[self.off_rs:self.off_rs + q].zero_()
        self.rs_parity = 0
        self.rs_dirty = False          # set by other users of the RS region (ring attention accumulators)
        torch.cuda.synchronize()
        dist.barrier(group)            # flag pages are zero everywhere before the first device barrier

    def take_epochs(self, slot: int, n: int = 1) -> int:
        """Epoch argument for the next launch on ``slot``: always 0 = "use the device counter" (the host-side count is
        kept only for debugging)."""
        self.epoch[slot] += n
        return 0

    def view(self, offset: int, shape, dtype) -> torch.Tensor:
        n = 1
        for s in shape:
            n *= s
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        return self.window[offset:offset + nbytes].view(dtype).view(*shape)

    def peer_ptr_array(self, offset: int):
        arr = (ctypes.c_void_p * self.world)(*[p + offset for p in self.peer_ptrs])
        return arr

    def barrier(self, slot: int = SLOT_BARRIER) -> None:
        N.call("stb_symm_barrier", self.flag_pages_dev, self.rank, self.world, slot, self.take_epochs(slot), N.stream_ptr())


def get_group(group, window_bytes: int | None = None) -> SymmGroup:
    """The cached ``SymmGroup`` (symmetric window, flag pages, multicast mapping) of a process group; allocated collectively on first use."""
    key = id(group)
    g = _groups.get(key)
    if g is None:
        g = SymmGroup(group, window_bytes or 512 * MB)
        _groups[key] = g
    return g


def reset() -> None:
    """Forget the cached symmetric groups (called when the process groups are torn down)."""
    _groups.clear()


# ------------------------------------------------------------------------------------------ ops
def all_gather_gemm(x_shard: torch.Tensor, weight: torch.Tensor, group, b_mn: bool = False, n_copy: int = 0,
                    flag_rows: int = 128):
    """``(y [T, N], x_full [T, K])`` with ``T = world · T_shard``; weight ``[N, K]`` (or ``[K, N]`` if ``b_mn``)."""
    g = get_group(group)
    if n_copy <= 0:
        # copy CTAs drive TMA bulk copies (~100 KB in flight each): two pairs hide a TP2 gather behind the GEMM, wider
        # groups move (world - 1) / world of the activation and get four
        n_copy = int(os.environ.get("SCALETORCH_B200_AG_COPY_CTAS", 4 if g.world <= 2 else 8))
    Ts, K = x_shard.shape
    T = Ts * g.world
    Nn = weight.shape[1] if b_mn else weight.shape[0]
    assert Ts * K * 2 <= g.region_bytes, "all-gather shard does not fit the symmetric window region"
    shard = g.view(g.off_ag, (Ts, K), torch.bfloat16)
    shard.copy_(x_shard)
    x_full = _empty((T, K), torch.bfloat16, x_shard.device)
    y = _empty((T, Nn), torch.bfloat16, x_shard.device)
    epoch = g.take_epochs(SLOT_AG, 2)
    N.call("stb_ag_gemm_bf16", x_full.data_ptr(), weight.data_ptr(), y.data_ptr(), T, Nn, K, weight.stride(0),
           y.stride(0), int(b_mn), g.bufs_dev, g.flag_pages_dev, g.off_ag, g.ready_flags.data_ptr(), epoch, flag_rows,
           n_copy, g.rank, g.world, SLOT_AG, N.stream_ptr())
    return y, x_full


def gemm_reduce_scatter(x: torch.Tensor, weight: torch.Tensor, group, b_mn: bool = False) -> torch.Tensor:
    """``reduce_scatter_rows(x · op(W)ᵀ)``: returns this rank's ``[T / world, N]`` row block of the summed product."""
    g = get_group(group)
    T, K = x.shape
    Nn = weight.shape[1] if b_mn else weight.shape[0]
    rows = T // g.world
    half = g.region_bytes // 2
    assert rows * Nn * 2 <= half, "reduce-scatter destination does not fit half of the symmetric window region"
    # Two destination buffers alternate.  The one this call adds into was zeroed on every rank during the previous call,
    # before that call's closing barrier; this call zeroes the other one for its successor.  One device barrier per call
    # (it makes the adds visible AND orders the zero-fill before the next call's adds) instead of zero + barrier + GEMM +
    # barrier.
    if g.rs_dirty:                                       # somebody else wrote into the region: restore the invariant
        g.window[g.off_rs:g.off_rs + g.region_bytes].zero_()
        g.barrier(SLOT_RS)
        g.rs_dirty = False
    off = g.off_rs + g.rs_parity * half
    other = g.off_rs + (1 - g.rs_parity) * half
    g.rs_parity ^= 1
    dest = g.view(off, (rows, Nn), torch.bfloat16)
    g.window[other:other + half].zero_()
    ptrs = g.peer_ptr_array(off)
    N.call("stb_gemm_rs_bf16", x.data_ptr(), weight.data_ptr(), ptrs, T, Nn, K, x.stride(0), weight.stride(0), Nn,
           int(b_mn), g.rank, g.world, N.stream_ptr())
    g.barrier(SLOT_RS)                                   # every rank's adds have landed; the other half is zero everywhere
    return dest.clone()                                  # the buffer is recycled two calls from now


def gemm_all_reduce(x: torch.Tensor, weight: torch.Tensor, group, b_mn: bool = False) -> torch.Tensor:
    """``all_reduce(x · op(W)ᵀ)``: the GEMM writes its partial into the window, then every rank sums the peers'
    partials straight out of their memory (one-shot) into a fresh local tensor."""
    from ..ops.gemm import gemm

    g = get_group(group)
    T = x.shape[0]
    Nn = weight.shape[1] if b_mn else weight.shape[0]
    assert T * Nn * 2 <= g.region_bytes
    part = g.view(g.off_ar, (T, Nn), torch.bfloat16)
    gemm(x, weight, b_mn=b_mn, out=part)
    return all_reduce_window_bf16(g, g.off_ar, T * Nn).view(T, Nn)


def all_reduce_window_bf16(g: SymmGroup, offset_bytes: int, n: int, sm_budget: int = 0) -> torch.Tensor:
    """Sum ``n`` bf16 elements at ``offset_bytes`` of every rank's window: in place through NVLS ``multimem`` when the group has a multicast mapping (returns a copy), otherwise one-shot peer loads into a fresh tensor."""
    if g.mc_ptr and n % 8 == 0 and not _no_multimem():
        # NVLS: in-switch reduction + broadcast in place on the window (each rank handles 1/world of the range)
        epoch = g.take_epochs(SLOT_AR, 2)
        N.call("stb_allreduce_bf16", g.bufs_dev, g.mc_ptr, g.flag_pages_dev, offset_bytes // 2, n, g.rank, g.world,
               SLOT_AR, epoch, sm_budget or 64, N.stream_ptr())
        return g.view(offset_bytes, (n,), torch.bfloat16).clone()
    out = _empty((n,), torch.bfloat16, g.device)
    epoch = g.take_epochs(SLOT_AR, 2)
    N.call("stb_allreduce_oneshot_bf16", g.bufs_dev, g.flag_pages_dev, offset_bytes // 2, out.data_ptr(), n, g.rank,
           g.world, SLOT_AR, epoch, sm_budget, N.stream_ptr())
    return out


def all_reduce_(t: torch.Tensor, group) -> torch.Tensor:
    """Sum ``t`` (bf16, contiguous) over the group through the window; returns a new tensor."""
    g = get_group(group)
    n = t.numel()
    assert n * 2 <= g.region_bytes and n % 8 == 0
    g.view(g.off_ar, (n,), torch.bfloat16).copy_(t.reshape(-1))
    return all_reduce_window_bf16(g, g.off_ar, n).view(t.shape)


class GradWindow:
    """fp32 gradient buffer allocated in symmetric memory so data-parallel buckets are reduced in place by
    ``stb_allreduce_f32`` (two-shot over peer memory, NVLS multimem when available) with the 1/world scale fused."""

    def __init__(self, group, numel: int):
        import torch.distributed._symmetric_memory as sm

        self.g = get_group(group)
        dev = self.g.device
        nbytes = (numel * 4 + (1 << 21) - 1) // (1 << 21) * (1 << 21)
        self.raw = sm.empty(nbytes, dtype=torch.uint8, device=dev)
        self.h = sm.rendezvous(self.raw, group=group)
        self.bufs_dev = int(self.h.buffer_ptrs_dev)
        self.mc_ptr = _multicast_ptr(self.h)
        import os

        # transport (profiles/dp_allreduce_sweep.md): NVLS multimem with 16 CTAs was the fastest setting at 2 and at 8
        # ranks once the step is replayed from a CUDA graph; SCALETORCH_B200_NO_MULTIMEM=1 selects the peer-load path
        if os.environ.get("SCALETORCH_B200_NO_MULTIMEM", "0") == "1":
            self.mc_ptr = 0
        if self.g.rank == 0 and os.environ.get("SCALETORCH_B200_VERBOSE", "0") == "1":
            print(f"[GradWindow] {nbytes >> 20} MiB, world {self.g.world}, multicast_ptr={'yes' if self.mc_ptr else 'no'}", flush=True)
        self.grad = self.raw[: numel * 4].view(torch.float32)
        self.grad.zero_()
        torch.cuda.synchronize()
        dist.barrier(group)

    def all_reduce_range_(self, start: int, numel: int, scale: float, sm_budget: int = 32) -> None:
        g = self.g
        epoch = g.take_epochs(SLOT_DP, 2)
        N.call("stb_allreduce_f32", self.bufs_dev, self.mc_ptr or None, g.flag_pages_dev, start, numel, float(scale),
               g.rank, g.world, SLOT_DP, epoch, sm_budget, N.stream_ptr())

    def reduce_scatter_range_(self, start: int, numel: int, scale: float, own_lo: int, own_hi: int,
                              sm_budget: int = 32) -> None:
        """Mean of ``[start, start + numel)`` over the group, kept only where it intersects this rank's shard
        ``[own_lo, own_hi)`` (``stb_reduce_scatter_f32``); every rank must call it (the kernel synchronises the group)."""
        g = self.g
        epoch = g.take_epochs(SLOT_DP, 2)
        N.call("stb_reduce_scatter_f32", self.bufs_dev, self.mc_ptr or None, g.flag_pages_dev, start, numel, own_lo,
               own_hi, float(scale), g.rank, g.world, SLOT_DP, epoch, sm_budget, N.stream_ptr())


class ParamWindow:
    """bf16 parameter buffer in symmetric memory: the sharded optimiser (``trainer/zero.py::PeerShardedAdamW``) updates
    its shard and writes the new values into every rank's copy from inside the AdamW kernel
    (``stb_adamw_shard_bcast``: ``multimem.st`` through the switch, or one store per peer).  Also holds 32 fp32 words per
    rank for the squared-norm partials (``stb_allreduce_tiny_f32``)."""

    def __init__(self, group, numel: int, dtype: torch.dtype):
        import torch.distributed._symmetric_memory as sm

        assert dtype == torch.bfloat16, "the peer-memory parameter window holds bf16 parameters"
        self.g = get_group(group)
        dev = self.g.device
        nbytes = (numel * 2 + (1 << 21) - 1) // (1 << 21) * (1 << 21)
        self.raw = sm.empty(nbytes + 4096, dtype=torch.uint8, device=dev)
        self.raw.zero_()
        self.h = sm.rendezvous(self.raw, group=group)
        self.bufs_dev = int(self.h.buffer_ptrs_dev)
        self.mc_ptr = 0 if _no_multimem() else _multicast_ptr(self.h)
        self.param = self.raw[: numel * 2].view(torch.bfloat16)
        self.scratch_off = nbytes // 4                        # fp32 element offset of the 32-word scratch
        self.scratch = self.raw[nbytes: nbytes + 128].view(torch.float32)
        torch.cuda.synchronize()
        dist.barrier(group)

    def all_reduce_scalars_(self, n: int) -> torch.Tensor:
        """Sums ``self.scratch[:n]`` over the group in place (n <= 32) and returns the view."""
        g = self.g
        N.call("stb_allreduce_tiny_f32", self.bufs_dev, g.flag_pages_dev, self.scratch_off, n, g.rank, g.world, SLOT_AR,
               g.take_epochs(SLOT_AR, 2), N.stream_ptr())
        return self.scratch[:n]


class EmbeddingGradExchange:
    """Data-parallel synchronisation of an embedding gradient by exchanging token rows (``stb_embedding_bwd_peers``).

    Installed on the embedding weight by ``DataParallelBucket`` (attribute ``_stb_sparse_sync``);
    ``ops.embedding.embedding_backward`` calls :meth:`accumulate` instead of scatter-adding locally.  Every
    micro-batch is exchanged, also under ``no_sync``: the contributions are identical on all ranks afterwards, so a
    later mean all-reduce of the same range (tied LM head) leaves them unchanged.
    """

    def __init__(self, group, scale: float, wait_for=None):
        self.g = get_group(group)
        self.scale = float(scale)
        self.wait_for = wait_for          # callable: make the current stream wait for the dense all-reduce of the range

    def fits(self, T: int, H: int) -> bool:
        return T * 8 + T * H * 2 + 4096 <= self.g.region_bytes

    def accumulate(self, flat_ids: torch.Tensor, dout2d: torch.Tensor, dst: torch.Tensor, vocab_start: int, vocab_local: int):
        g = self.g
        T, H = dout2d.shape
        ids_off = g.off_ag
        dout_off = g.off_ag + (T * 8 + 4095) // 4096 * 4096
        g.view(ids_off, (T,), torch.int64).copy_(flat_ids)
        g.view(dout_off, (T, H), torch.bfloat16).copy_(dout2d)
        if self.wait_for is not None:
            self.wait_for()
        N.call("stb_embedding_bwd_peers", g.bufs_dev, g.flag_pages_dev, ids_off, dout_off, dst.data_ptr(), T, H,
               int(vocab_start), int(vocab_local), self.scale, g.rank, g.world, SLOT_EMB, g.take_epochs(SLOT_EMB, 2), 32,
               N.stream_ptr())
