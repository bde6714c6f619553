"""Context parallelism: ring attention over contiguous sequence shards.

Reference: ``scaletorch/parallel/context_parallel/context_parallel.py:54-514``.  Same interface
(``apply_context_parallel``, ``ring_attention``, ``update_rope_for_context_parallel``,
``update_output_and_lse``); the per-block math is the flash kernel returning ``(out, lse)`` instead
of an eager materialised ``[B,h,S,S]`` score matrix, K/V travel un-expanded (``Hkv`` heads, not ``Hq``),
and the block backward consumes the global ``(out, lse)`` so no per-block recomputation of the
normaliser is needed.  Layout is ``[B, S_local, H, D]``.

Causality follows the reference: rank ``r`` attends to blocks ``j ≤ r`` (diagonal block masked).
"""
from __future__ import annotations

import os

import torch

from ... import ops
from ...env import ENV_CONTEXT_PARALLEL
from ..process_group import process_group_manager as pgm
from .cp_comms import ContextCommunicator


def apply_context_parallel(model: torch.nn.Module) -> torch.nn.Module:
    """Reference-compatible switch (``context_parallel.py:54-80``).  Models of this framework pick ring
    attention from their ``ParallelContext`` at construction, so this only records the env var for
    tools that look at it and re-resolves the backend of already-built attention modules."""
    os.environ[ENV_CONTEXT_PARALLEL] = "1" if (pgm and pgm.cp_world_size > 1) else "0"
    if pgm and pgm.cp_world_size > 1:
        for m in model.modules():
            if hasattr(m, "_attn_backend_name"):
                m._attn_backend_name = "ring"
    return model


def update_output_and_lse(out, lse, block_out, block_lse):
    """Numerically stable merge of two partial attention results (reference ``:367-424``).

    out ``[B,S,H,D]`` (any float dtype, merged in fp32), lse ``[B,H,S]`` fp32.
    """
    if out is None:
        return block_out.float(), block_lse
    new_lse = torch.logaddexp(lse, block_lse)
    w_old = torch.exp(lse - new_lse).transpose(1, 2).unsqueeze(-1)      # [B,S,H,1]
    w_new = torch.exp(block_lse - new_lse).transpose(1, 2).unsqueeze(-1)
    return out * w_old + block_out.float() * w_new, new_lse


def update_rows(out, lse, block_out, block_lse, q_lo: int):
    """:func:`update_output_and_lse` restricted to query rows ``[q_lo:]`` (zig-zag partial blocks)."""
    if q_lo == 0:
        return update_output_and_lse(out, lse, block_out, block_lse)
    o2, l2 = update_output_and_lse(out[:, q_lo:], lse[:, :, q_lo:], block_out, block_lse)
    out[:, q_lo:] = o2
    lse[:, :, q_lo:] = l2
    return out, lse


class RingAttentionFunc(torch.autograd.Function):
    """Ring attention on c10d p2p: K/V blocks rotate around the CP ring while each rank runs block attention and merges ``(out, lse)``; the backward rotates K/V again together with the dK/dV accumulators.  Contiguous or zig-zag chunks; the CPU / fallback counterpart of ``fused_ring``."""
    @staticmethod
    def forward(ctx, q, k, v, sm_scale, is_causal, group, send_rank, recv_rank, cp_rank, cp_size):
        from . import zigzag

        zz = zigzag.enabled() and is_causal
        S = q.shape[1]
        comm = ContextCommunicator("kv", group, send_rank, recv_rank)
        k_cur, v_cur = k.contiguous(), v.contiguous()
        out, lse = None, None
        for step in range(cp_size):
            if step + 1 != cp_size:
                k_next = comm.send_recv(k_cur)
                v_next = comm.send_recv(v_cur)
                comm.commit()
            plan = zigzag.block_plan(step, cp_rank, cp_size, S, is_causal, zz)
            if plan is not None:
                q_lo, k_hi, cf = plan
                b_out, b_lse = ops.attention.attention_block_fwd(q[:, q_lo:], k_cur[:, :k_hi], v_cur[:, :k_hi], cf, sm_scale)
                out, lse = update_rows(out, lse, b_out, b_lse, q_lo)
            if step + 1 != cp_size:
                comm.wait()
                k_cur, v_cur = k_next, v_next
        out = out.to(q.dtype)
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.cfg = (sm_scale, is_causal, group, send_rank, recv_rank, cp_rank, cp_size, zz)
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse = ctx.saved_tensors
        from . import zigzag

        sm_scale, is_causal, group, send_rank, recv_rank, cp_rank, cp_size, zz = ctx.cfg
        S = q.shape[1]
        kv_comm = ContextCommunicator("kv-bwd", group, send_rank, recv_rank)
        d_comm = ContextCommunicator("dkv", group, send_rank, recv_rank)
        dout = dout.contiguous()
        dq = torch.zeros_like(q, dtype=torch.float32)
        k_cur, v_cur = k.contiguous(), v.contiguous()
        dk_acc = dv_acc = None          # gradient of the block currently held, travelling with it
        for step in range(cp_size):
            if step + 1 != cp_size:
                k_next = kv_comm.send_recv(k_cur)
                v_next = kv_comm.send_recv(v_cur)
                kv_comm.commit()
            plan = zigzag.block_plan(step, cp_rank, cp_size, S, is_causal, zz)
            if plan is not None:
                q_lo, k_hi, cf = plan
                if q_lo == 0 and k_hi == S:
                    b_dq, b_dk, b_dv = ops.attention.attention_block_bwd(dout, q, k_cur, v_cur, out, lse, cf, sm_scale)
                    dq += b_dq.float()
                    b_dk, b_dv = b_dk.float(), b_dv.float()
                else:
                    p_dq, p_dk, p_dv = ops.attention.attention_block_bwd(
                        dout[:, q_lo:].contiguous(), q[:, q_lo:], k_cur[:, :k_hi], v_cur[:, :k_hi],
                        out[:, q_lo:].contiguous(), lse[:, :, q_lo:].contiguous(), cf, sm_scale)
                    dq[:, q_lo:] += p_dq.float()
                    b_dk = torch.zeros_like(k_cur, dtype=torch.float32)
                    b_dv = torch.zeros_like(v_cur, dtype=torch.float32)
                    b_dk[:, :k_hi] = p_dk.float()
                    b_dv[:, :k_hi] = p_dv.float()
            else:
                b_dk = torch.zeros_like(k_cur, dtype=torch.float32)
                b_dv = torch.zeros_like(v_cur, dtype=torch.float32)
            if step > 0:
                d_comm.wait()           # dk_acc/dv_acc for this block arrived from the previous rank
                dk_acc, dv_acc = dk_recv + b_dk, dv_recv + b_dv
            else:
                dk_acc, dv_acc = b_dk, b_dv
            # pass the accumulated gradient along with the block (one extra hop at the end brings it home)
            dk_recv = d_comm.send_recv(dk_acc)
            dv_recv = d_comm.send_recv(dv_acc)
            d_comm.commit()
            if step + 1 != cp_size:
                kv_comm.wait()
                k_cur, v_cur = k_next, v_next
        d_comm.wait()
        return dq.to(q.dtype), dk_recv.to(k.dtype), dv_recv.to(v.dtype), None, None, None, None, None, None, None


def ring_attention(q, k, v, sm_scale: float, is_causal: bool = True, group=None) -> torch.Tensor:
    """q ``[B,S_l,Hq,D]``, k/v ``[B,S_l,Hkv,D]`` (this rank's contiguous sequence chunk) → ``[B,S_l,Hq,D]``."""
    if group is None:
        if not pgm or pgm.cp_world_size == 1:
            return ops.flash_attention(q, k, v, is_causal, sm_scale)[0]
        group, send_rank, recv_rank = pgm.cp_group, pgm.cp_send_rank, pgm.cp_recv_rank
        cp_rank, cp_size = pgm.cp_rank, pgm.cp_world_size
        if q.is_cuda and os.environ.get("SCALETORCH_B200_FUSED_COMM", "1") == "1":
            from . import fused_ring

            if fused_ring.available(q, k, group):   # K/V pulled from the peers' HBM by the attention kernel itself
                return fused_ring.fused_ring_attention(q, k, v, sm_scale, is_causal, group)
    else:
        import torch.distributed as dist

        cp_rank, cp_size = dist.get_rank(group), dist.get_world_size(group)
        ranks = dist.get_process_group_ranks(group)
        send_rank, recv_rank = ranks[(cp_rank + 1) % cp_size], ranks[(cp_rank - 1) % cp_size]
    return RingAttentionFunc.apply(q, k, v, sm_scale, is_causal, group, send_rank, recv_rank, cp_rank, cp_size)


def update_rope_for_context_parallel(cos: torch.Tensor, sin: torch.Tensor, cp_rank: int | None = None,
                                     cp_size: int | None = None):
    """Slices ``[S, D]`` RoPE tables to this rank's contiguous chunk (reference ``:427-473``).  The
    framework's own models index the global tables with a position offset instead."""
    if cp_rank is None:
        if not pgm:
            return cos, sin
        cp_rank, cp_size = pgm.cp_rank, pgm.cp_world_size
    S = cos.shape[0]
    if S % cp_size != 0:
        raise ValueError(f"sequence length {S} must be divisible by cp size {cp_size}")
    n = S // cp_size
    return cos[cp_rank * n:(cp_rank + 1) * n], sin[cp_rank * n:(cp_rank + 1) * n]
