"""RMSNorm (optionally fused with the residual add) and the fused per-head
QK-RMSNorm + rotate-half RoPE.

Reference semantics: ``RMSNorm`` in ``scaletorch/models/attention_utils.py:247-271``
(fp32 statistics, normalised value cast to the activation dtype *before* the weight
multiply), ``apply_rotary_pos_emb`` / ``get_cos_sin`` in ``attention_utils.py:170-239``
and the q_norm/k_norm → RoPE sequence of ``scaletorch/models/model_qwen3.py:205-217``.
The RoPE tables are kept in fp32 (the reference stores them in bf16).
"""
from __future__ import annotations

import torch

from .. import _native as N

_scratch: dict[tuple, torch.Tensor] = {}


def _get_scratch(key: tuple, numel: int, device, dtype=torch.float32) -> torch.Tensor:
    buf = _scratch.get(key)
    if buf is None or buf.numel() < numel:
        buf = torch.empty(max(numel, 1), dtype=dtype, device=device)
        _scratch[key] = buf
    return buf


def _param_grad_out(param: torch.Tensor):
    """Returns (fp32 destination, accumulate flag, is_main_grad)."""
    mg = getattr(param, "main_grad", None)
    if mg is not None:
        return mg.view(param.shape), bool(getattr(param, "grad_written", False)), True
    return torch.empty(param.shape, dtype=torch.float32, device=param.device), False, False


def _finish_param_grad(param, dst, is_main):
    if is_main:
        param.grad_written = True
        hook = getattr(param, "grad_ready_hook", None)
        if hook is not None:
            hook()
        return None
    return dst.to(param.dtype)


# --------------------------------------------------------------------------- oracle
def rmsnorm_ref(x: torch.Tensor, w: torch.Tensor, eps: float) -> torch.Tensor:
    """PyTorch oracle: ``w * (x * rsqrt(mean(x^2) + eps))`` with the statistics in fp32 and the cast before the weight multiply."""
    xf = x.float()
    xn = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype)
    return w * xn


def rotate_half(x: torch.Tensor) -> torch.Tensor:
    """``cat(-x2, x1)`` for ``x = cat(x1, x2)`` along the last dimension (the rotate-half RoPE convention)."""
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


def get_cos_sin(seq_len: int, head_dim: int, base: float = 500000.0, device="cpu") -> tuple[torch.Tensor, torch.Tensor]:
    """fp32 ``[seq_len, head_dim]`` tables, halves duplicated (rotate-half convention)."""
    assert head_dim % 2 == 0
    inv = 1.0 / (base ** (torch.arange(0, head_dim, 2, dtype=torch.float64, device="cpu") / head_dim))
    pos = torch.arange(seq_len, dtype=torch.float64, device="cpu")
    ang = torch.outer(pos, inv)
    cos = torch.cos(ang).repeat(1, 2).float().to(device)
    sin = torch.sin(ang).repeat(1, 2).float().to(device)
    return cos, sin


def apply_rotary_pos_emb(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """x: ``[..., S, D]``; cos/sin broadcastable ``[S, D]``.  Computed in fp32."""
    xf = x.float()
    return (xf * cos.float() + rotate_half(xf) * sin.float()).to(x.dtype)


def qknorm_rope_ref(qkv, wq, wk, cos, sin, n_q, n_kv, head_dim, seq_len, pos_offset, eps, positions=None):
    """qkv ``[T, (Hq+2Hkv)*D]`` → rotated ``[T, (Hq+Hkv)*D]`` (oracle)."""
    T = qkv.shape[0]
    nqk = n_q + n_kv
    x = qkv[:, : nqk * head_dim].reshape(T, nqk, head_dim)
    if wq is not None:
        w = torch.cat([wq.expand(n_q, head_dim), wk.expand(n_kv, head_dim)], 0)
        xf = x.float()
        xn = (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype)
        x = w * xn
    if positions is None:
        positions = torch.arange(T, device=qkv.device) % seq_len + pos_offset
    c = cos[positions.long()].unsqueeze(1)
    s = sin[positions.long()].unsqueeze(1)
    xf = x.float()
    out = (xf * c + rotate_half(xf) * s).to(qkv.dtype)
    return out.reshape(T, nqk * head_dim)


# --------------------------------------------------------------------------- RMSNorm
class _AddRMSNorm(torch.autograd.Function):
    """(x, res?) -> (h = x + res, y = rmsnorm(h) * w)."""

    @staticmethod
    def forward(ctx, x, res, w, eps):
        H = x.shape[-1]
        x2 = x.reshape(-1, H)
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        if res is not None:
            r2 = res.reshape(-1, H)
            h = torch.empty_like(x2)
        else:
            r2, h = None, x2
        N.call("stb_rmsnorm_fwd", x2.data_ptr(), N.ptr(r2), w.data_ptr(), h.data_ptr() if res is not None else None,
               y.data_ptr(), rstd.data_ptr(), rows, H, float(eps), N.stream_ptr())
        ctx.save_for_backward(h, w, rstd)
        ctx.has_res = res is not None
        ctx.shape = x.shape
        if res is not None:
            return h.view(x.shape), y.view(x.shape)
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, *grads):
        h, w, rstd = ctx.saved_tensors
        rows, H = h.shape
        if ctx.has_res:
            dh, dy = grads
        else:
            dh, dy = None, grads[0]
        if dy is None:  # y unused: only the residual stream carries gradient
            return dh, dh, None, None
        dy2 = dy.reshape(rows, H)
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dres = None
        if dh is not None:
            dres = dh.reshape(rows, H)
            if not dres.is_contiguous():
                dres = dres.contiguous()
        dx = torch.empty_like(h)
        dwdst, acc, is_main = _param_grad_out(w)
        nscr = N.call_long("stb_rmsnorm_bwd_scratch_elems", rows, H)
        scratch = _get_scratch(("rms", h.device.index), nscr, h.device)
        N.call("stb_rmsnorm_bwd", dy2.data_ptr(), h.data_ptr(), w.data_ptr(), rstd.data_ptr(), N.ptr(dres),
               dx.data_ptr(), dwdst.data_ptr(), int(acc), scratch.data_ptr(), rows, H, N.stream_ptr())
        dw = _finish_param_grad(w, dwdst, is_main)
        dx = dx.view(ctx.shape)
        if ctx.has_res:
            return dx, dx, dw, None
        return dx, None, dw, None


def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    """RMSNorm over the last dimension: the fused kernel for bf16 CUDA tensors, the oracle otherwise."""
    if N.use_native(x) and x.dtype == torch.bfloat16:
        return _AddRMSNorm.apply(x.contiguous(), None, w, eps)
    return rmsnorm_ref(x, w, eps)


def add_rmsnorm(x: torch.Tensor, res: torch.Tensor, w: torch.Tensor, eps: float = 1e-6):
    """Returns ``(h, y)`` with ``h = x + res`` (new residual stream) and ``y = rmsnorm(h)``."""
    if N.use_native(x) and x.dtype == torch.bfloat16:
        return _AddRMSNorm.apply(x.contiguous(), res.contiguous(), w, eps)
    h = x + res
    return h, rmsnorm_ref(h, w, eps)


# ------------------------------------------------------------------ QK-norm + RoPE
class _QKNormRope(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, wq, wk, cos, sin, n_q, n_kv, head_dim, seq_len, pos_offset, eps, positions):
        T = qkv.shape[0]
        use_norm = wq is not None
        out = torch.empty((T, (n_q + n_kv) * head_dim), dtype=qkv.dtype, device=qkv.device)
        rstd = torch.empty((T * (n_q + n_kv),), dtype=torch.float32, device=qkv.device) if use_norm else None
        N.call("stb_qknorm_rope_fwd", qkv.data_ptr(), N.ptr(wq), N.ptr(wk), cos.data_ptr(), sin.data_ptr(),
               N.ptr(positions), out.data_ptr(), N.ptr(rstd), T, seq_len, n_q, n_kv, head_dim, pos_offset,
               int(use_norm), float(eps), N.stream_ptr())
        ctx.save_for_backward(qkv, wq, wk, cos, sin, rstd, positions)
        ctx.cfg = (n_q, n_kv, head_dim, seq_len, pos_offset, use_norm)
        return out

    @staticmethod
    def backward(ctx, dqk):
        qkv, wq, wk, cos, sin, rstd, positions = ctx.saved_tensors
        n_q, n_kv, head_dim, seq_len, pos_offset, use_norm = ctx.cfg
        T = qkv.shape[0]
        dqk = dqk.contiguous()
        # d(qkv): q,k parts written by the kernel, the v part is zero here (the fused attention op in
        # ops/attention.py writes dV into the same buffer instead of going through this Function).
        dqkv = torch.zeros_like(qkv)
        dwq = dwk = None
        if use_norm:
            dq_dst, _, q_main = _param_grad_out(wq)
            dk_dst, _, k_main = _param_grad_out(wk)
            # the kernel accumulates with atomics: start from zero unless accumulating into main_grad
            if not (q_main and getattr(wq, "grad_written", False)):
                dq_dst.zero_()
            if not (k_main and getattr(wk, "grad_written", False)):
                dk_dst.zero_()
        N.call("stb_qknorm_rope_bwd", dqk.data_ptr(), qkv.data_ptr(), N.ptr(wq), N.ptr(wk), cos.data_ptr(),
               sin.data_ptr(), N.ptr(positions), N.ptr(rstd), dqkv.data_ptr(),
               dq_dst.data_ptr() if use_norm else None, dk_dst.data_ptr() if use_norm else None,
               T, seq_len, n_q, n_kv, head_dim, pos_offset, int(use_norm), N.stream_ptr())
        if use_norm:
            dwq = _finish_param_grad(wq, dq_dst, q_main)
            dwk = _finish_param_grad(wk, dk_dst, k_main)
        return dqkv, dwq, dwk, None, None, None, None, None, None, None, None, None


def qknorm_rope(qkv, wq, wk, cos, sin, n_q, n_kv, head_dim, seq_len, pos_offset=0, eps=1e-6, positions=None):
    """Fused per-head RMSNorm (if ``wq``/``wk`` given) + RoPE on the q and k heads of a packed
    ``[T, (Hq+2Hkv)*D]`` projection.  Returns ``[T, (Hq+Hkv)*D]``; V stays in ``qkv``.
    """
    if N.use_native(qkv) and qkv.dtype == torch.bfloat16 and head_dim == 128:
        pos32 = positions.to(torch.int32) if positions is not None else None
        return _QKNormRope.apply(qkv, wq, wk, cos, sin, n_q, n_kv, head_dim, seq_len, pos_offset, eps, pos32)
    return qknorm_rope_ref(qkv, wq, wk, cos, sin, n_q, n_kv, head_dim, seq_len, pos_offset, eps, positions)
