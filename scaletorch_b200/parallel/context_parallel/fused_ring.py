"""Ring attention without a ring: every rank publishes its K/V block in symmetric memory and every other rank pulls
the blocks it needs straight out of that peer's HBM over NVLink while it is busy with its own (diagonal) block.

Forward: K/V are copied once into the CP group's window; after one on-device barrier a side stream pulls the needed
peer blocks into local staging buffers (one pass over NVLink per block -- measured: letting the attention kernel's
TMA read the peer window directly re-reads each block once per query tile and makes the kernel NVLink-bound, 178 ms
vs 121 ms per step on Qwen3-1.7B cp2), step 0 (own block, causal) runs meanwhile, then step ``s`` runs the tcgen05
flash kernel on the staged block of rank ``(r - s) mod cp`` and folds the partial result into ``(out, lse)`` in its
epilogue (``stb_flash_attn_fwd_block``).  Backward: each rank runs the block backward kernel per block it attended to;
dQ accumulates in fp32 inside the kernel and the (dK, dV) contributions are added straight into the fp32 accumulators
in the window of the rank that owns the block (``red.global.add`` over NVLink), so after one barrier every owner holds
its complete gradient.  No send/recv, no forwarding hops, no eager math between the attention launches, K/V travel
un-expanded (``Hkv`` heads).  The staging copies run on the copy engines (peer ``cudaMemcpyAsync`` on a side stream):
they take no SMs from the attention kernels they overlap with.

Reference: ``RingAttentionFunc`` with batched isend/irecv and eager per-block attention
(``scaletorch/parallel/context_parallel/context_parallel.py:111-263``, ``cp_comms.py:72-222``).
"""
from __future__ import annotations

import torch

from ... import ops
from ...runtime import fused_ops as F
from ...runtime import symm
from . import zigzag

SLOT_CP = 5
_pull_streams: dict[int, torch.cuda.Stream] = {}


def _pull_stream(device: torch.device) -> torch.cuda.Stream:
    idx = device.index if device.index is not None else torch.cuda.current_device()
    st = _pull_streams.get(idx)
    if st is None:
        st = _pull_streams[idx] = torch.cuda.Stream(device=device)
    return st


def _stage_peer_blocks(g, steps, shape, n):
    """Start pulling the K/V blocks of the ranks in ``steps`` (``{step: src}``) into local buffers on the side
    stream; returns ``{step: (k, v, event)}``.  The caller waits on the event before using a block."""
    if not steps:
        return {}
    cur = torch.cuda.current_stream()
    side = _pull_stream(g.device)
    published = torch.cuda.Event()
    published.record(cur)
    staged = {}
    for step in steps:
        staged[step] = (torch.empty(shape, dtype=torch.bfloat16, device=g.device),
                        torch.empty(shape, dtype=torch.bfloat16, device=g.device), torch.cuda.Event())
    with torch.cuda.stream(side):
        side.wait_event(published)
        for step, src in steps.items():
            ks, vs, ev = staged[step]
            ks.copy_(g.window_h.get_buffer(src, shape, torch.bfloat16, g.off_spare // 2))
            vs.copy_(g.window_h.get_buffer(src, shape, torch.bfloat16, (g.off_spare + n * 2) // 2))
            ev.record(side)
    return staged


def available(q: torch.Tensor, k: torch.Tensor, group) -> bool:
    """True when the peer-memory ring applies: bf16 CUDA tensors with head_dim 128, the block kernels present, symmetric memory set up for the group and large enough for K/V and the fp32 dK/dV accumulators."""
    if not (q.is_cuda and q.dtype == torch.bfloat16 and q.shape[-1] == 128 and ops.native_attention_available()):
        return False
    if not symm.group_ready(group) or not symm.has_kernel("stb_flash_attn_bwd_block"):
        return False
    g = F.get_group(group)
    kv_bytes = k.numel() * 2
    # window: [K | V] bf16 in the spare region; fp32 [dK | dV] accumulators of the own block in the rs + ar regions
    return 2 * kv_bytes <= g.region_bytes and 2 * k.numel() * 4 <= 2 * g.region_bytes


class FusedRingAttention(torch.autograd.Function):
    """No eager math between the attention launches: the (out, lse) merge happens in the forward kernel's epilogue
    (fp32 running output), dQ accumulates in fp32 inside the backward kernel, and every rank adds its (dK, dV)
    contributions straight into the block owner's fp32 accumulators over NVLink (``red.global.add``)."""

    @staticmethod
    def forward(ctx, q, k, v, sm_scale, is_causal, group):
        g = F.get_group(group)
        r, cp = g.rank, g.world
        B, S, Hkv, D = k.shape
        n = k.numel()
        kw = g.view(g.off_spare, (B, S, Hkv, D), torch.bfloat16)
        vw = g.view(g.off_spare + n * 2, (B, S, Hkv, D), torch.bfloat16)
        kw.copy_(k)
        vw.copy_(v)
        g.barrier(SLOT_CP)                      # every rank's K/V block is published
        # contiguous split: blocks in the future of every local query are skipped under a causal mask; zig-zag split:
        # every peer block contributes half a block (zigzag.block_plan)
        zz = zigzag.enabled() and is_causal
        plans = {s_: zigzag.block_plan(s_, r, cp, S, is_causal, zz) for s_ in range(1, cp)}
        steps = {s_: (r - s_) % cp for s_ in range(1, cp) if plans[s_] is not None}
        staged = _stage_peer_blocks(g, steps, (B, S, Hkv, D), n)
        out = F._empty(q.shape, q.dtype, q.device)
        lse = torch.empty((B, q.shape[2], S), dtype=torch.float32, device=q.device)
        acc = torch.empty(q.shape, dtype=torch.float32, device=q.device) if steps else None
        ops.attention.ring_block_fwd_(q, k, v, out, lse, acc, False, 0, is_causal, sm_scale)
        for step in steps:
            kb, vb, ev = staged[step]
            torch.cuda.current_stream().wait_event(ev)
            q_lo, k_hi, cf = plans[step]
            ops.attention.ring_block_fwd_(q[:, q_lo:], kb[:, :k_hi], vb[:, :k_hi], out, lse, acc, True, q_lo, cf, sm_scale)
        g.barrier(SLOT_CP)                      # every pull has completed: windows may be overwritten
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.cfg = (sm_scale, is_causal, group, zz)
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse = ctx.saved_tensors
        sm_scale, is_causal, group, zz = ctx.cfg
        g = F.get_group(group)
        r, cp = g.rank, g.world
        B, S, Hkv, D = k.shape
        n = k.numel()
        dout = dout.contiguous()
        kw = g.view(g.off_spare, (B, S, Hkv, D), torch.bfloat16)
        vw = g.view(g.off_spare + n * 2, (B, S, Hkv, D), torch.bfloat16)
        kw.copy_(k)
        vw.copy_(v)
        # fp32 accumulators of the gradient of MY block, in my window: every rank that attended to the block adds into them
        off_dk, off_dv = g.off_rs, g.off_rs + n * 4
        dk32 = g.view(off_dk, (B, S, Hkv, D), torch.float32)
        dv32 = g.view(off_dv, (B, S, Hkv, D), torch.float32)
        g.view(off_dk, (2 * n,), torch.float32).zero_()
        g.rs_dirty = True
        g.barrier(SLOT_CP)                      # K/V published, accumulators zeroed
        plans = {s_: zigzag.block_plan(s_, r, cp, S, is_causal, zz) for s_ in range(1, cp)}
        steps = {s_: (r - s_) % cp for s_ in range(1, cp) if plans[s_] is not None}
        staged = _stage_peer_blocks(g, steps, (B, S, Hkv, D), n)
        dq32 = torch.empty(q.shape, dtype=torch.float32, device=q.device)
        delta = torch.empty((B, q.shape[2], S), dtype=torch.float32, device=q.device)
        kv_strides = (S * Hkv * D, Hkv * D, D)
        for step in [0] + list(steps):
            src = (r - step) % cp
            if step == 0:
                kb, vb = k, v
                q_lo, k_hi, cf = 0, S, is_causal
            else:
                kb, vb, ev = staged[step]
                torch.cuda.current_stream().wait_event(ev)
                q_lo, k_hi, cf = plans[step]
            ops.attention.ring_block_bwd_(dout, q[:, q_lo:], kb[:, :k_hi], vb[:, :k_hi], out, lse, delta, dq32,
                                          g.peer_ptrs[src] + off_dk, g.peer_ptrs[src] + off_dv, kv_strides, q_lo, cf,
                                          sm_scale, first=(step == 0))
        g.barrier(SLOT_CP)                      # every rank's contributions have landed in the owners' accumulators
        return dq32.to(q.dtype), dk32.to(k.dtype), dv32.to(v.dtype), None, None, None


def fused_ring_attention(q, k, v, sm_scale: float, is_causal: bool, group) -> torch.Tensor:
    """Context-parallel attention over peer memory (see the module docstring); q / k / v are this rank's ``[B, S_local, H, D]`` chunks, the result is its chunk of the full-sequence attention."""
    return FusedRingAttention.apply(q, k.contiguous(), v.contiguous(), sm_scale, is_causal, group)
