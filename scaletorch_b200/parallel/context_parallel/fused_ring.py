"""Ring attention without a ring: every rank publishes its K/V block in symmetric memory and the flash-attention
kernel of every other rank TMA-loads the blocks it needs straight out of that peer's HBM over NVLink.

Forward: K/V are copied once into the CP group's window; after one on-device barrier, step ``s`` runs the tcgen05
flash kernel with K/V tensor maps built on rank ``(r - s) mod cp``'s window (step 0 = own block, causal) and the
partial ``(out, lse)`` results are merged.  Backward: each rank computes the (dK, dV) contributions of the blocks
it attended to into its own window (one slot per source block); after a barrier the owner of each block sums the
contributions of the ranks that used it by reading their windows.  No send/recv, no forwarding hops, K/V travel
un-expanded (``Hkv`` heads).

Reference: ``RingAttentionFunc`` with batched isend/irecv and eager per-block attention
(``scaletorch/parallel/context_parallel/context_parallel.py:111-263``, ``cp_comms.py:72-222``).
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ... import ops
from ...runtime import fused_ops as F
from ...runtime import symm
from .context_parallel import update_output_and_lse

SLOT_CP = 5


def available(q: torch.Tensor, k: torch.Tensor, group) -> bool:
    if not (q.is_cuda and q.dtype == torch.bfloat16 and q.shape[-1] == 128 and ops.native_attention_available()):
        return False
    if not symm.group_ready(group):
        return False
    g = F.get_group(group)
    kv_bytes = k.numel() * 2
    # window: [K | V] forward + cp x [dK | dV] fp32 backward slots must fit the spare + rs regions
    return 2 * kv_bytes <= g.region_bytes and g.world * 2 * k.numel() * 4 <= 2 * g.region_bytes


class FusedRingAttention(torch.autograd.Function):
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
        out, lse = None, None
        for step in range(cp):
            src = (r - step) % cp
            if is_causal and step > r:
                continue                        # block lies in the future of every local query
            if step == 0:
                kb, vb = kw, vw
            else:                               # tensors aliasing the peer's window (TMA reads go over NVLink)
                kb = g.window_h.get_buffer(src, (B, S, Hkv, D), torch.bfloat16, g.off_spare // 2)
                vb = g.window_h.get_buffer(src, (B, S, Hkv, D), torch.bfloat16, (g.off_spare + n * 2) // 2)
            b_out, b_lse = ops.attention.attention_block_fwd(q, kb, vb, is_causal and step == 0, sm_scale)
            out, lse = update_output_and_lse(out, lse, b_out, b_lse)
        g.barrier(SLOT_CP)                      # nobody still reads my block when I overwrite it next time
        out = out.to(q.dtype)
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.cfg = (sm_scale, is_causal, group)
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse = ctx.saved_tensors
        sm_scale, is_causal, group = ctx.cfg
        g = F.get_group(group)
        r, cp = g.rank, g.world
        B, S, Hkv, D = k.shape
        n = k.numel()
        dout = dout.contiguous()
        kw = g.view(g.off_spare, (B, S, Hkv, D), torch.bfloat16)
        vw = g.view(g.off_spare + n * 2, (B, S, Hkv, D), torch.bfloat16)
        kw.copy_(k)
        vw.copy_(v)
        # contribution slots: slot j holds this rank's (dK, dV) for block j, bf16, in the rs+ar regions
        slot_bytes = 2 * n * 2
        base = g.off_rs

        def slot(j):
            dk = g.view(base + j * slot_bytes, (B, S, Hkv, D), torch.bfloat16)
            dv = g.view(base + j * slot_bytes + n * 2, (B, S, Hkv, D), torch.bfloat16)
            return dk, dv

        g.barrier(SLOT_CP)
        dq = torch.zeros(q.shape, dtype=torch.float32, device=q.device)
        for step in range(cp):
            src = (r - step) % cp
            if is_causal and step > r:
                continue
            if step == 0:
                kb, vb = kw, vw
            else:
                kb = g.window_h.get_buffer(src, (B, S, Hkv, D), torch.bfloat16, g.off_spare // 2)
                vb = g.window_h.get_buffer(src, (B, S, Hkv, D), torch.bfloat16, (g.off_spare + n * 2) // 2)
            dkj, dvj = slot(src)
            dqj = torch.empty_like(q)
            ops.attention._flash_bwd(dout, q, kb, vb, out, lse, dqj, dkj, dvj, is_causal and step == 0, sm_scale)
            dq += dqj.float()
        g.barrier(SLOT_CP)                      # all contributions are written
        # block r was used by ranks r, r+1, ..., cp-1 (causal) or by everyone
        dk = torch.zeros(k.shape, dtype=torch.float32, device=k.device)
        dv = torch.zeros(v.shape, dtype=torch.float32, device=v.device)
        users = range(r, cp) if is_causal else range(cp)
        for u in users:
            off = base + r * slot_bytes
            if u == r:
                dku, dvu = slot(r)
            else:
                dku = g.window_h.get_buffer(u, (B, S, Hkv, D), torch.bfloat16, off // 2)
                dvu = g.window_h.get_buffer(u, (B, S, Hkv, D), torch.bfloat16, (off + n * 2) // 2)
            dk += dku.float()
            dv += dvu.float()
        g.barrier(SLOT_CP)                      # contribution slots may be reused
        return dq.to(q.dtype), dk.to(k.dtype), dv.to(v.dtype), None, None, None


def fused_ring_attention(q, k, v, sm_scale: float, is_causal: bool, group) -> torch.Tensor:
    return FusedRingAttention.apply(q, k.contiguous(), v.contiguous(), sm_scale, is_causal, group)
