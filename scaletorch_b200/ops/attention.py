"""Causal flash attention with native GQA, and the fused
``qkv → (QK-norm) → RoPE → attention`` block used by the decoder layers.

Reference: ``flash_attention`` = ``F.scaled_dot_product_attention`` after an explicit GQA
expansion of K/V (``scaletorch/models/attention_utils.py:130-152``,
``scaletorch/models/model_qwen3.py:219-238``).  Here K/V are never expanded: the kernel maps
``Hq/Hkv`` query heads onto each KV head.  Tensors use the token-major ``[B, S, H, D]`` layout the
QKV GEMM produces, so no transposes are materialised either.

``flash_attention`` returns ``(out, lse)``; ``lse`` (natural-log sum-exp, ``[B, H, S]`` fp32) is what
ring attention needs to merge partial results (``scaletorch/parallel/context_parallel/
context_parallel.py:367-424``).
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from .. import _native as N
from . import norm as _norm


# --------------------------------------------------------------------------- oracle
def attention_ref(q, k, v, causal=True, scale=None, q_offset: int = 0):
    """fp32 eager attention. q ``[B,Sq,Hq,D]``, k/v ``[B,Sk,Hkv,D]`` → (out ``[B,Sq,Hq,D]``, lse ``[B,Hq,Sq]``).

    ``q_offset``: global position of q[0] relative to k[0] (for ring / chunked causal masking).
    """
    B, Sq, Hq, D = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    g = Hq // Hkv
    scale = scale if scale is not None else 1.0 / math.sqrt(D)
    qf = q.float().permute(0, 2, 1, 3).reshape(B, Hkv, g, Sq, D)
    kf = k.float().permute(0, 2, 1, 3)
    vf = v.float().permute(0, 2, 1, 3)
    s = torch.einsum("bhgqd,bhkd->bhgqk", qf, kf) * scale
    if causal:
        qi = torch.arange(Sq, device=q.device)[:, None] + q_offset
        ki = torch.arange(Sk, device=q.device)[None, :]
        s = s.masked_fill(ki > qi, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)
    p = torch.exp(s - lse.unsqueeze(-1))
    p = torch.nan_to_num(p, nan=0.0)
    o = torch.einsum("bhgqk,bhkd->bhgqd", p, vf)
    out = o.reshape(B, Hq, Sq, D).permute(0, 2, 1, 3).to(q.dtype)
    return out, lse.reshape(B, Hq, Sq)


def native_attention_fwd_available() -> bool:
    """True when the in-tree library exports the flash-attention forward kernel."""
    lib = N.load()
    return lib is not None and hasattr(lib, "stb_flash_attn_fwd")


def native_attention_available() -> bool:
    """Forward *and* backward kernels present (what training needs)."""
    lib = N.load()
    return lib is not None and hasattr(lib, "stb_flash_attn_fwd") and hasattr(lib, "stb_flash_attn_bwd")


def _flash_fwd(q, k, v, causal, scale):
    """q/k/v: [B,S,H,D] (last dim contiguous, arbitrary token/head strides)."""
    B, Sq, Hq, D = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    out = torch.empty((B, Sq, Hq, D), dtype=q.dtype, device=q.device)
    lse = torch.empty((B, Hq, Sq), dtype=torch.float32, device=q.device)
    N.call("stb_flash_attn_fwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), lse.data_ptr(),
           B, Sq, Sk, Hq, Hkv, D,
           q.stride(0), q.stride(1), q.stride(2), k.stride(0), k.stride(1), k.stride(2),
           v.stride(0), v.stride(1), v.stride(2), float(scale), int(causal), N.stream_ptr())
    return out, lse


def _flash_bwd(do, q, k, v, out, lse, dq, dk, dv, causal, scale):
    B, Sq, Hq, D = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    delta = torch.empty((B, Hq, Sq), dtype=torch.float32, device=q.device)
    N.call("stb_flash_attn_bwd", do.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(),
           lse.data_ptr(), delta.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(),
           B, Sq, Sk, Hq, Hkv, D,
           q.stride(0), q.stride(1), q.stride(2), k.stride(0), k.stride(1), k.stride(2),
           v.stride(0), v.stride(1), v.stride(2),
           dq.stride(0), dq.stride(1), dq.stride(2), dk.stride(0), dk.stride(1), dk.stride(2),
           dv.stride(0), dv.stride(1), dv.stride(2), float(scale), int(causal), N.stream_ptr())


_warned_head_dim: set = set()


def fused_bwd_available() -> bool:
    """One-kernel backward (``csrc/attention_bwd_fused.cu``): S / dP and the elementwise pass run once, dQ is reduce-added
    into an fp32 buffer.  ``SCALETORCH_B200_ATTN_BWD_FUSED=0`` selects the two-kernel backward."""
    import os

    lib = N.load()
    return (lib is not None and hasattr(lib, "stb_flash_attn_bwd_fused")
            and os.environ.get("SCALETORCH_B200_ATTN_BWD_FUSED", "1") == "1")


def _flash_bwd_fused(do, q, k, v, out, lse, dq32, dk, dv, causal, scale):
    """``dq32``: zero-initialised fp32 ``[B,Sq,Hq,D]`` (strided view allowed), accumulated into; ``dk`` / ``dv`` bf16."""
    B, Sq, Hq, D = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    delta = torch.empty((B, Hq, Sq), dtype=torch.float32, device=q.device)
    N.call("stb_attn_bwd_delta", do.data_ptr(), out.data_ptr(), delta.data_ptr(), B, Sq, Hq, N.stream_ptr())
    N.call("stb_flash_attn_bwd_fused", do.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), lse.data_ptr(),
           delta.data_ptr(), dq32.data_ptr(), dk.data_ptr(), dv.data_ptr(), B, Sq, Sk, Hq, Hkv, D,
           q.stride(0), q.stride(1), q.stride(2), k.stride(0), k.stride(1), k.stride(2),
           v.stride(0), v.stride(1), v.stride(2),
           dq32.stride(0), dq32.stride(1), dq32.stride(2), dk.stride(0), dk.stride(1), dk.stride(2),
           dv.stride(0), dv.stride(1), dv.stride(2), float(scale), int(causal), N.stream_ptr())


def _sdpa(q, k, v, causal, scale):
    """Library attention for head dims the native kernel does not cover (it is written for D = 128, the head dim of
    every Qwen3 / Llama-2/3 size the reference benchmarks).  Returns ``(out, lse)`` like the native path -- ``lse`` comes
    from the flash backend so ring attention keeps working -- and says so once per head dim instead of silently
    changing kernels."""
    D = q.shape[-1]
    if D not in _warned_head_dim:
        _warned_head_dim.add(D)
        import warnings

        warnings.warn(f"flash_attention: head_dim {D} runs on the PyTorch SDPA flash backend; the native tcgen05 kernel "
                      f"covers head_dim 128 only")
    g = q.shape[2] // k.shape[2]
    qt, kt, vt = q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)
    if g > 1:
        kt, vt = kt.repeat_interleave(g, dim=1), vt.repeat_interleave(g, dim=1)
    try:
        res = torch.ops.aten._scaled_dot_product_flash_attention(qt, kt, vt, 0.0, causal, False, scale=scale)
        return res[0].transpose(1, 2), res[1].detach()
    except Exception:  # backend refuses this head dim / dtype: plain SDPA, lse from the fp32 oracle only if asked for
        o = F.scaled_dot_product_attention(qt, kt, vt, is_causal=causal, scale=scale)
        return o.transpose(1, 2), None


class _FlashAttention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, causal, scale):
        out, lse = _flash_fwd(q, k, v, causal, scale)
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.cfg = (causal, scale)
        ctx.mark_non_differentiable(lse)
        return out, lse

    @staticmethod
    def backward(ctx, do, _dlse):
        q, k, v, out, lse = ctx.saved_tensors
        causal, scale = ctx.cfg
        dk, dv = torch.empty_like(k), torch.empty_like(v)
        if fused_bwd_available():
            dq32 = torch.zeros(q.shape, dtype=torch.float32, device=q.device)
            _flash_bwd_fused(do.contiguous(), q, k, v, out, lse, dq32, dk, dv, causal, scale)
            return dq32.to(q.dtype), dk, dv, None, None
        dq = torch.empty_like(q)
        _flash_bwd(do.contiguous(), q, k, v, out, lse, dq, dk, dv, causal, scale)
        return dq, dk, dv, None, None


def flash_attention(q, k, v, causal: bool = True, scale: float | None = None):
    """q ``[B,Sq,Hq,D]``, k/v ``[B,Sk,Hkv,D]`` → ``(out [B,Sq,Hq,D], lse [B,Hq,Sq])``."""
    scale = scale if scale is not None else 1.0 / math.sqrt(q.shape[-1])
    if N.use_native(q) and q.dtype == torch.bfloat16:
        if native_attention_available() and q.shape[-1] == 128:
            return _FlashAttention.apply(q, k, v, causal, scale)
        return _sdpa(q, k, v, causal, scale)
    return attention_ref(q, k, v, causal, scale)


# ------------------------------------------------- fused qkv -> norm/rope -> attention
class _QKVAttention(torch.autograd.Function):
    """One autograd node for ``qkv → QK-norm → RoPE → causal GQA attention``.

    The backward writes dQ/dK/dV straight into one packed ``dqkv`` buffer (the attention kernel
    takes output strides) and runs the norm/RoPE backward in place on its q,k columns, so the
    gradient of the packed projection is produced without concatenations or zero-fills.
    """

    @staticmethod
    def forward(ctx, qkv, wq, wk, cos, sin, B, S, n_q, n_kv, D, pos_offset, eps, causal, scale):
        T = B * S
        use_norm = wq is not None
        nqk = n_q + n_kv
        qk = torch.empty((T, nqk * D), dtype=qkv.dtype, device=qkv.device)
        rstd = torch.empty((T * nqk,), dtype=torch.float32, device=qkv.device) if use_norm else None
        N.call("stb_qknorm_rope_fwd", qkv.data_ptr(), N.ptr(wq), N.ptr(wk), cos.data_ptr(), sin.data_ptr(),
               None, qk.data_ptr(), N.ptr(rstd), T, S, n_q, n_kv, D, pos_offset, int(use_norm), float(eps),
               N.stream_ptr())
        qk4 = qk.view(B, S, nqk, D)
        q, k = qk4[:, :, :n_q], qk4[:, :, n_q:]
        v = qkv.view(B, S, n_q + 2 * n_kv, D)[:, :, nqk:]
        out, lse = _flash_fwd(q, k, v, causal, scale)
        ctx.save_for_backward(qkv, qk, wq, wk, cos, sin, rstd, out, lse)
        ctx.cfg = (B, S, n_q, n_kv, D, pos_offset, use_norm, causal, scale)
        return out.view(B, S, n_q * D)

    @staticmethod
    def backward(ctx, do):
        qkv, qk, wq, wk, cos, sin, rstd, out, lse = ctx.saved_tensors
        B, S, n_q, n_kv, D, pos_offset, use_norm, causal, scale = ctx.cfg
        T = B * S
        nqk = n_q + n_kv
        qk4 = qk.view(B, S, nqk, D)
        q, k = qk4[:, :, :n_q], qk4[:, :, n_q:]
        v = qkv.view(B, S, n_q + 2 * n_kv, D)[:, :, nqk:]
        dqkv = torch.empty_like(qkv)
        d4 = dqkv.view(B, S, n_q + 2 * n_kv, D)
        dqk = torch.empty_like(qk)
        dqk4 = dqk.view(B, S, nqk, D)
        do4 = do.reshape(B, S, n_q, D)
        if not do4.is_contiguous():
            do4 = do4.contiguous()
        dq32 = None
        if fused_bwd_available():
            # dQ accumulates in fp32 (TMA reduce-add from the fused kernel) and is consumed as fp32 by the norm / RoPE
            # backward below: no conversion pass, no bf16 round trip
            dq32 = torch.zeros((B, S, n_q, D), dtype=torch.float32, device=qkv.device)
            _flash_bwd_fused(do4, q, k, v, out, lse, dq32, dqk4[:, :, n_q:], d4[:, :, nqk:], causal, scale)
        else:
            _flash_bwd(do4, q, k, v, out, lse, dqk4[:, :, :n_q], dqk4[:, :, n_q:], d4[:, :, nqk:], causal, scale)
        dwq = dwk = None
        if use_norm:
            dq_dst, _, q_main = _norm._param_grad_out(wq)
            dk_dst, _, k_main = _norm._param_grad_out(wk)
            if not (q_main and getattr(wq, "grad_written", False)):
                dq_dst.zero_()
            if not (k_main and getattr(wk, "grad_written", False)):
                dk_dst.zero_()
        if dq32 is not None:
            N.call("stb_qknorm_rope_bwd_f32q", dqk.data_ptr(), dq32.data_ptr(), qkv.data_ptr(), N.ptr(wq), N.ptr(wk),
                   cos.data_ptr(), sin.data_ptr(), None, N.ptr(rstd), dqkv.data_ptr(),
                   dq_dst.data_ptr() if use_norm else None, dk_dst.data_ptr() if use_norm else None,
                   T, S, n_q, n_kv, D, pos_offset, int(use_norm), N.stream_ptr())
        else:
            N.call("stb_qknorm_rope_bwd", dqk.data_ptr(), qkv.data_ptr(), N.ptr(wq), N.ptr(wk), cos.data_ptr(),
                   sin.data_ptr(), None, N.ptr(rstd), dqkv.data_ptr(),
                   dq_dst.data_ptr() if use_norm else None, dk_dst.data_ptr() if use_norm else None,
                   T, S, n_q, n_kv, D, pos_offset, int(use_norm), N.stream_ptr())
        if use_norm:
            dwq = _norm._finish_param_grad(wq, dq_dst, q_main)
            dwk = _norm._finish_param_grad(wk, dk_dst, k_main)
        return (dqkv, dwq, dwk) + (None,) * 11


def qkv_attention(qkv, wq, wk, cos, sin, B, S, n_q, n_kv, head_dim, pos_offset=0, eps=1e-6, causal=True,
                  scale=None, attn_fn=None, positions=None):
    """``qkv [B*S, (Hq+2Hkv)*D]`` → attention output ``[B, S, Hq*D]``.

    ``attn_fn(q, k, v, causal, scale) -> out`` overrides the local attention (ring attention under
    context parallelism); in that case the norm/RoPE and attention run as separate autograd nodes.
    """
    scale = scale if scale is not None else 1.0 / math.sqrt(head_dim)
    if (attn_fn is None and positions is None and N.use_native(qkv) and qkv.dtype == torch.bfloat16 and head_dim == 128
            and native_attention_available()):
        return _QKVAttention.apply(qkv, wq, wk, cos, sin, B, S, n_q, n_kv, head_dim, pos_offset, eps, causal, scale)
    nqk = n_q + n_kv
    qk = _norm.qknorm_rope(qkv, wq, wk, cos, sin, n_q, n_kv, head_dim, S, pos_offset, eps, positions)
    qk4 = qk.view(B, S, nqk, head_dim)
    q, k = qk4[:, :, :n_q], qk4[:, :, n_q:]
    v = qkv.view(B, S, n_q + 2 * n_kv, head_dim)[:, :, nqk:]
    if attn_fn is not None:
        out = attn_fn(q, k, v, causal, scale)
    else:
        out = flash_attention(q, k, v, causal, scale)[0]
    return out.reshape(B, S, n_q * head_dim)


# ------------------------------------------------------- block-level API (ring attention)
def attention_block_bwd_ref(do, q, k, v, out, lse, causal, scale, q_offset: int = 0):
    """fp32 block backward given the *global* ``out`` / ``lse`` of the query rows (oracle).

    Returns ``(dq, dk, dv)`` contributions of this K/V block: ``P = exp(S - lse)``,
    ``dS = P ∘ (dO·Vᵀ − rowsum(dO ∘ O))``.
    """
    B, Sq, Hq, D = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    g = Hq // Hkv
    qf = q.float().permute(0, 2, 1, 3).reshape(B, Hkv, g, Sq, D)
    kf = k.float().permute(0, 2, 1, 3)
    vf = v.float().permute(0, 2, 1, 3)
    dof = do.float().permute(0, 2, 1, 3).reshape(B, Hkv, g, Sq, D)
    of = out.float().permute(0, 2, 1, 3).reshape(B, Hkv, g, Sq, D)
    s = torch.einsum("bhgqd,bhkd->bhgqk", qf, kf) * scale
    if causal:
        qi = torch.arange(Sq, device=q.device)[:, None] + q_offset
        ki = torch.arange(Sk, device=q.device)[None, :]
        s = s.masked_fill(ki > qi, float("-inf"))
    p = torch.exp(s - lse.reshape(B, Hkv, g, Sq, 1))
    delta = (dof * of).sum(-1, keepdim=True)
    dp = torch.einsum("bhgqd,bhkd->bhgqk", dof, vf)
    ds = p * (dp - delta) * scale
    dq = torch.einsum("bhgqk,bhkd->bhgqd", ds, kf).reshape(B, Hq, Sq, D).permute(0, 2, 1, 3)
    dk = torch.einsum("bhgqk,bhgqd->bhkd", ds, qf).permute(0, 2, 1, 3)
    dv = torch.einsum("bhgqk,bhgqd->bhkd", p, dof).permute(0, 2, 1, 3)
    return dq.to(q.dtype), dk.to(k.dtype), dv.to(v.dtype)


def ring_block_fwd_(q, k, v, out, lse, acc, merge: bool, q_lo: int, causal: bool, scale: float) -> None:
    """One key block of a blockwise attention, merged in the kernel's epilogue (``stb_flash_attn_fwd_block``).

    ``q [B,Sq,Hq,D]`` are query rows ``[q_lo, q_lo + Sq)`` of ``out [B,S,Hq,D]`` (bf16), ``acc`` (fp32 running output,
    optional) and ``lse [B,Hq,S]``; with ``merge`` the block is folded into the result over the blocks seen so far
    (reference: the eager ``update_out_and_lse``, ``context_parallel.py:367-424``)."""
    B, Sq, Hq, D = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    N.call("stb_flash_attn_fwd_block", q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(), lse.data_ptr(),
           N.ptr(acc), int(merge), int(q_lo), out.shape[1], B, Sq, Sk, Hq, Hkv, D,
           q.stride(0), q.stride(1), q.stride(2), k.stride(0), k.stride(1), k.stride(2),
           v.stride(0), v.stride(1), v.stride(2), float(scale), int(causal), N.stream_ptr())


def ring_block_bwd_(do, q, k, v, out, lse, delta, dq32, dk32_ptr: int, dv32_ptr: int, kv_strides, q_lo: int, causal: bool,
                    scale: float, first: bool) -> None:
    """One key block of the backward of a blockwise attention (``stb_flash_attn_bwd_block``): ``dq32`` (fp32, full
    ``[B,S,Hq,D]``) is written on the ``first`` block and accumulated afterwards; this rank's (dK, dV) contributions are
    ``red.add``-ed into the fp32 accumulators at ``dk32_ptr`` / ``dv32_ptr`` -- raw device addresses, usually in the memory
    of the GPU that owns the block -- laid out with ``kv_strides`` (batch, token, head; in elements)."""
    B, Sq, Hq, D = q.shape
    Sk, Hkv = k.shape[1], k.shape[2]
    dq = dq32[:, q_lo:]
    N.call("stb_flash_attn_bwd_block", do.data_ptr(), q.data_ptr(), k.data_ptr(), v.data_ptr(), out.data_ptr(),
           lse.data_ptr(), delta.data_ptr(), dq.data_ptr(), dk32_ptr, dv32_ptr, B, Sq, Sk, Hq, Hkv, D,
           q.stride(0), q.stride(1), q.stride(2), k.stride(0), k.stride(1), k.stride(2),
           v.stride(0), v.stride(1), v.stride(2), dq.stride(0), dq.stride(1), dq.stride(2),
           kv_strides[0], kv_strides[1], kv_strides[2], kv_strides[0], kv_strides[1], kv_strides[2],
           float(scale), int(causal), int(q_lo), out.shape[1], int(first), int(not first), N.stream_ptr())


def attention_block_fwd(q, k, v, causal, scale):
    """(out, lse) of one K/V block -- native flash kernel on CUDA, oracle elsewhere."""
    if N.use_native(q) and q.dtype == torch.bfloat16 and native_attention_available() and q.shape[-1] == 128:
        return _flash_fwd(q, k, v, causal, scale)
    return attention_ref(q, k, v, causal, scale)


def attention_block_bwd(do, q, k, v, out, lse, causal, scale):
    """(dq, dk, dv) of one K/V block given the global out / lse."""
    if N.use_native(q) and q.dtype == torch.bfloat16 and native_attention_available() and q.shape[-1] == 128:
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        _flash_bwd(do.contiguous(), q, k, v, out, lse, dq, dk, dv, causal, scale)
        return dq, dk, dv
    return attention_block_bwd_ref(do, q, k, v, out, lse, causal, scale)
