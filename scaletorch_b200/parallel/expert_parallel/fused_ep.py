"""Expert-parallel dispatch → grouped expert MLP → combine over NVLink peer memory (``csrc/ep_dispatch.cu``).

What the c10d path (``ep_comms.py``; reference ``scaletorch/parallel/expert_parallel/ep_comms.py:30-171``) does
with 3-5 ``all_to_all_single`` calls, two host synchronisations for the split sizes and a local re-sort, this path
does with three kernels and no host synchronisation:

1. ``stb_ep_exchange_counts``: the per-expert pair counts of every rank become visible to every rank (a [ep, E]
   table, one device barrier).  From it each *sender* computes, with a handful of tensor ops, the exact row every one
   of its (token, k) pairs occupies in the destination rank's receive region -- which is laid out expert-major with
   each expert's segment padded to 128 rows, i.e. directly consumable by the ragged grouped GEMM.
2. ``stb_ep_push``: token rows are stored straight into the peers' regions.
3. grouped GEMMs on the region; the down-projection writes its output into the symmetric *return* region.
4. ``stb_ep_pull``: every token pulls its k expert rows from their owners, applies the gate weights and sums.

The backward is the mirror image (the pull's backward is a scaled push of the output gradient, the push's backward
is an unweighted pull of the input gradients), so gradients reach the hidden states and the router.

The receive capacity is static (``capacity_factor × T·k`` rows, never more than the worst case ``ep × T·k``): pairs
that do not fit are dropped on the sender (their contribution is zero in both directions) and counted in
``FusedEPState.dropped``; with ``capacity_factor >= ep`` the path is exactly dropless.
"""
from __future__ import annotations

import ctypes
import os

import torch
import torch.distributed as dist

from ... import _native as N
from ...ops.activation import swiglu
from ...ops.grouped import expert_counts
from ...ops.grouped_native import GroupedLayout, _GroupedLinearPadded
from ...runtime import fused_ops

_P, _I, _L, _U = ctypes.c_void_p, ctypes.c_int, ctypes.c_long, ctypes.c_uint
N.register("stb_ep_exchange_counts", [_P, _P, _L, _P, _I, _P, _I, _I, _I, _U, _P])
N.register("stb_ep_push", [_P, _P, _L, _P, _P, _P, _P, _L, _I, _I, _I, _I, _I, _U, _I, _P])
N.register("stb_ep_pull", [_P, _P, _L, _P, _P, _P, _P, _P, _L, _I, _I, _I, _I, _I, _U, _I, _P])
N.register("stb_ep_plan", [_P, _P, _P, _L, _I, _I, _I, _I, _P, _P, _P, _P, _P, _P, _I, _P])
N.register("stb_ep_zero_padding", [_P, _P, _P, _I, _I, _I, _P])
N.register("stb_copy_rows_dyn", [_P, _P, _P, _L, _I, _P])

SLOT_EP = 6
_windows: dict = {}


def default_capacity_factor() -> float | None:
    """``None`` = dropless (receive regions sized for the worst case ``ep × T·k`` rows).  A smaller factor bounds the
    regions and every row-padded activation at ``factor × T·k`` rows and DROPS the pairs that do not fit -- an explicit
    opt-in (``--ep_capacity_factor`` / ``SCALETORCH_B200_EP_CAPACITY``), never the default."""
    v = os.environ.get("SCALETORCH_B200_EP_CAPACITY")
    return float(v) if v else None


class EPWindow:
    """Symmetric receive (X) / return (Y) regions + count table of one EP group, sized for ``rows_cap × H`` bf16."""

    def __init__(self, group, rows_cap: int, hidden: int, num_experts: int):
        import torch.distributed._symmetric_memory as sm

        self.g = fused_ops.get_group(group, 64 * fused_ops.MB)
        dev = self.g.device
        self.rows_cap, self.hidden = rows_cap, hidden
        region = rows_cap * hidden * 2
        region = (region + 4095) // 4096 * 4096
        table = (self.g.world * num_experts * 4 + 4095) // 4096 * 4096
        self.off_x, self.off_y, self.off_table = 0, region, 2 * region
        self.raw = sm.empty(2 * region + table, dtype=torch.uint8, device=dev)
        self.raw.zero_()
        self.h = sm.rendezvous(self.raw, group=group)
        self.bufs_dev = int(self.h.buffer_ptrs_dev)
        self.x = self.raw[: rows_cap * hidden * 2].view(torch.bfloat16).view(rows_cap, hidden)
        self.y = self.raw[region: region + rows_cap * hidden * 2].view(torch.bfloat16).view(rows_cap, hidden)
        torch.cuda.synchronize()
        dist.barrier(group)


def get_window(group, rows_cap: int, hidden: int, num_experts: int) -> EPWindow:
    """The cached symmetric receive / return window of an EP group for this capacity and shape (allocated collectively on first use)."""
    key = (id(group), rows_cap, hidden, num_experts)
    w = _windows.get(key)
    if w is None:
        w = EPWindow(group, rows_cap, hidden, num_experts)
        _windows[key] = w
    return w


def reset() -> None:
    """Forget the cached EP windows (called when the process groups are torn down)."""
    _windows.clear()


def ep_plan(table: torch.Tensor, flat: torch.Tensor, counts: torch.Tensor, rank: int, e_local: int, rows_cap: int):
    """Destination of every ``(token, k)`` pair of rank ``rank`` -- pure tensor math, no communication.  The GPU path
    computes the same plan with ONE kernel (``stb_ep_plan``, ``csrc/moe_router.cu``); this function is its oracle
    (``tests/test_ep_plan_cpu.py``, ``tests/test_kernels_gpu.py``) with positions in the sender's stable order.

    ``table [ep, E]``: pairs every source rank sends to every global expert (identical on all ranks after the count
    exchange); ``flat [P]``: this rank's expert ids in pair order; ``counts [E]`` = ``table[rank]``.  The receive region
    of every rank is expert-major with each expert's segment padded to 128 rows, and inside a segment the rows are
    ordered by source rank, then by the sender's stable order.  Returns ``(row [P] int32 (−1 = over capacity),
    dest_rank [P] int32, dropped (scalar), offs_pad [e_local + 1] of *this* rank's experts (clamped to the capacity),
    tokens_per_expert [e_local] of this rank's experts)``.
    """
    ep = table.shape[0]
    P = flat.numel()
    tot = table.sum(0)
    pad = ((tot + 127) // 128 * 128).view(ep, e_local)
    offs_pad_all = (torch.cumsum(pad, 1) - pad).reshape(-1)                  # start of every expert on its owner
    prefix = (torch.cumsum(table, 0) - table)[rank]                          # rows of lower source ranks
    offs_local = torch.cumsum(counts, 0) - counts
    sort_idx = torch.argsort(flat, stable=True)
    inv = torch.empty_like(sort_idx)
    inv[sort_idx] = torch.arange(P, device=flat.device)
    pos = offs_pad_all[flat] + prefix[flat] + inv - offs_local[flat]
    over = pos >= rows_cap
    row = torch.where(over, torch.full_like(pos, -1), pos).to(torch.int32).contiguous()
    dest = torch.div(flat, e_local, rounding_mode="floor").to(torch.int32).contiguous()
    zero = torch.zeros(1, dtype=torch.int64, device=flat.device)
    offs_pad = torch.cat([zero, torch.cumsum(pad[rank], 0)]).clamp_(max=rows_cap)
    return row, dest, over.sum(), offs_pad, tot.view(ep, e_local)[rank]


class FusedEPState:
    """Routing plan of one MoE layer call (all tensors on the device, built by two kernels: the count exchange and
    ``stb_ep_plan``)."""

    def __init__(self, win: EPWindow, idx: torch.Tensor, num_experts: int, plan=None):
        g = win.g
        ep, rank = g.world, g.rank
        T, k = idx.shape
        e_local = num_experts // ep
        P = T * k
        dev = idx.device
        if plan is not None and plan[0].dtype == torch.int32 and idx.dtype == torch.int32:
            pos, counts = plan                               # from the routing kernel (ops/moe_routing.py)
            flat = idx.reshape(-1)
        else:                                                # ids from elsewhere: positions by a stable sort
            flat64 = idx.reshape(-1).long()
            c64 = expert_counts(flat64, num_experts)
            order = torch.argsort(flat64, stable=True)
            inv = torch.empty_like(order)
            inv[order] = torch.arange(P, device=dev)
            pos = (inv - (torch.cumsum(c64, 0) - c64)[flat64]).to(torch.int32)
            counts, flat = c64.to(torch.int32), flat64.to(torch.int32)
        table = torch.empty((ep, num_experts), dtype=torch.int32, device=dev)
        N.call("stb_ep_exchange_counts", win.bufs_dev, g.flag_pages_dev, win.off_table, counts.data_ptr(), num_experts,
               table.data_ptr(), rank, ep, SLOT_EP, g.take_epochs(SLOT_EP, 1), N.stream_ptr())
        n_tiles = win.rows_cap // 128
        meta = torch.zeros(1 + (e_local + 1) + e_local + n_tiles, dtype=torch.int32, device=dev)
        self.dropped = meta[0:1]
        offs_pad = meta[1:e_local + 2]
        self.tokens_per_expert = meta[e_local + 2:2 * e_local + 2]
        tile_expert = meta[2 * e_local + 2:]
        self.row = torch.empty(P, dtype=torch.int32, device=dev)
        self.rank_of = torch.empty(P, dtype=torch.int32, device=dev)
        N.call("stb_ep_plan", table.data_ptr(), flat.data_ptr(), pos.data_ptr(), P, num_experts, ep, rank, win.rows_cap,
               self.row.data_ptr(), self.rank_of.data_ptr(), self.dropped.data_ptr(), offs_pad.data_ptr(),
               self.tokens_per_expert.data_ptr(), tile_expert.data_ptr(), n_tiles, N.stream_ptr())
        self.win, self.T, self.k, self.P, self.e_local = win, T, k, P, e_local
        self.offs_pad = offs_pad
        self.layout = GroupedLayout.from_device_tables(offs_pad, tile_expert, win.rows_cap, e_local)

    # -- raw kernels ---------------------------------------------------------------------------------------
    def push(self, src: torch.Tensor, scale: torch.Tensor | None) -> None:
        """rows of ``src [T, H]`` → peers' X regions; only the (< 128 per expert) padding rows are zeroed first."""
        win, g = self.win, self.win.g
        N.call("stb_ep_zero_padding", win.x.data_ptr(), self.offs_pad.data_ptr(), self.tokens_per_expert.data_ptr(),
               self.e_local, win.hidden, win.rows_cap, N.stream_ptr())
        N.call("stb_ep_push", win.bufs_dev, g.flag_pages_dev, win.off_x, src.data_ptr(), self.rank_of.data_ptr(),
               self.row.data_ptr(), scale.data_ptr() if scale is not None else None, self.P, self.k, win.hidden,
               g.rank, g.world, SLOT_EP, g.take_epochs(SLOT_EP, 2), 0, N.stream_ptr())

    def pull(self, weights: torch.Tensor | None, keep_rows: bool):
        """Σ_k w · (row of the owners' Y regions) → ``[T, H]`` (and the raw rows ``[T·k, H]`` if asked)."""
        win, g = self.win, self.win.g
        out = torch.empty((self.T, win.hidden), dtype=torch.bfloat16, device=win.x.device)
        rows = torch.empty((self.P, win.hidden), dtype=torch.bfloat16, device=win.x.device) if keep_rows else None
        N.call("stb_ep_pull", win.bufs_dev, g.flag_pages_dev, win.off_y, out.data_ptr(),
               rows.data_ptr() if rows is not None else None, self.rank_of.data_ptr(), self.row.data_ptr(),
               weights.data_ptr() if weights is not None else None, self.T, self.k, win.hidden, g.rank, g.world,
               SLOT_EP, g.take_epochs(SLOT_EP, 2), 0, N.stream_ptr())
        return out, rows


class _Dispatch(torch.autograd.Function):
    """x ``[T, H]`` → my experts' padded input rows ``[rows_cap, H]`` (a private copy: it is saved for wgrad)."""

    @staticmethod
    def forward(ctx, x, st: FusedEPState):
        ctx.st = st
        st.push(x.contiguous(), None)
        # private copy of the rows in use (saved for the expert wgrad; the window is recycled by the next layer)
        xp = torch.empty_like(st.win.x)
        N.call("stb_copy_rows_dyn", st.win.x.data_ptr(), xp.data_ptr(), st.offs_pad[st.e_local:].data_ptr(),
               st.win.hidden * 2, st.win.rows_cap, N.stream_ptr())
        return xp

    @staticmethod
    def backward(ctx, d_xpad):
        st = ctx.st
        if d_xpad.data_ptr() != st.win.y.data_ptr():
            st.win.y.copy_(d_xpad)
        dx, _ = st.pull(None, keep_rows=False)
        return dx, None


class _Combine(torch.autograd.Function):
    """expert outputs in the Y region + gate weights ``[T, k]`` → ``[T, H]``."""

    @staticmethod
    def forward(ctx, y_pad, weights, st: FusedEPState):
        if y_pad.data_ptr() != st.win.y.data_ptr():
            st.win.y.copy_(y_pad)
        w = weights.reshape(-1).float().contiguous()
        out, rows = st.pull(w, keep_rows=True)
        ctx.st = st
        ctx.save_for_backward(rows, w)
        ctx.w_dtype = weights.dtype
        return out

    @staticmethod
    def backward(ctx, dout):
        st = ctx.st
        rows, w = ctx.saved_tensors
        dout = dout.contiguous()
        st.push(dout, w)                                     # d y_pad = w · dout[token] lands in the owners' X regions
        d_ypad = st.win.x                                     # consumed by the grouped dgrad/wgrad right away
        dw = torch.bmm(rows.view(st.T, st.k, -1), dout.unsqueeze(-1)).view(st.T, st.k).to(ctx.w_dtype)
        return d_ypad, dw, None


def fused_moe_ep(x: torch.Tensor, weights: torch.Tensor, idx: torch.Tensor, gate_up_proj: torch.Tensor,
                 down_proj: torch.Tensor, num_experts: int, group, capacity_factor: float | None = None, plan=None):
    """Full EP MoE branch: returns ``(out [T, H], FusedEPState)``.  ``plan``: ``(pos, counts)`` of the routing kernel."""
    T, H = x.shape
    k = idx.shape[1]
    ep = dist.get_world_size(group)
    e_local = num_experts // ep
    cf = default_capacity_factor() if capacity_factor is None else capacity_factor
    cf = float(ep) if cf is None else min(float(cf), float(ep))      # dropless unless a capacity was asked for
    rows_cap = (int(cf * T * k) + e_local * 127 + 127) // 128 * 128
    win = get_window(group, rows_cap, H, num_experts)
    st = FusedEPState(win, idx, num_experts, plan)
    xp = _Dispatch.apply(x, st)
    gu = _GroupedLinearPadded.apply(xp, gate_up_proj, st.layout, {"dx": win.y})
    y = _GroupedLinearPadded.apply(swiglu(gu), down_proj, st.layout, {"out": win.y})
    return _Combine.apply(y, weights, st), st


def available(group, x: torch.Tensor, gate_up_proj: torch.Tensor, down_proj: torch.Tensor) -> bool:
    """True when the peer-memory EP path applies: bf16 CUDA activations, the push / pull and grouped-GEMM kernels present, expert shapes the grouped GEMM supports, symmetric memory ready for the group."""
    from ...ops.grouped_native import supported
    from ...runtime import symm

    return (x.is_cuda and x.dtype == torch.bfloat16 and symm.has_kernel("stb_ep_push")
            and symm.has_kernel("stb_grouped_gemm_bf16") and supported(gate_up_proj) and supported(down_proj)
            and x.shape[1] % 8 == 0 and symm.group_ready(group))
