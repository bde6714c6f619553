"""Fused AdamW and global-norm gradient clipping over flat buffers.

Reference: ``torch.optim.AdamW(fused=True)`` chosen in
``scaletorch/trainer/model_builder.py:119-134`` and ``clip_grad_norm_`` in
``scaletorch/trainer/train_step.py:122-136``.  Because parameters, gradients and optimiser
state live in contiguous flat buffers (``runtime/flat_params.py``), the whole optimiser step is
three launches: sum-of-squares → clip coefficient (device scalar, never read on the host) →
AdamW with the coefficient folded in.
"""
from __future__ import annotations

import torch

from .. import _native as N

_scratch: dict[int, torch.Tensor] = {}


def _sumsq_scratch(device: torch.device) -> torch.Tensor:
    key = device.index or 0
    buf = _scratch.get(key)
    if buf is None:
        nbytes = N.call_long("stb_sumsq_scratch_bytes")
        buf = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


def sum_of_squares(g: torch.Tensor) -> torch.Tensor:
    """fp32 scalar tensor (on device) = Σ g²."""
    if N.use_native(g) and g.dtype == torch.float32:
        out = torch.empty(1, dtype=torch.float32, device=g.device)
        N.call("stb_sumsq", g.data_ptr(), g.numel(), _sumsq_scratch(g.device).data_ptr(), out.data_ptr(), N.stream_ptr())
        return out
    return g.float().pow(2).sum().reshape(1)


def clip_coef_from_sumsq(sumsq: torch.Tensor, max_norm: float) -> tuple[torch.Tensor, torch.Tensor]:
    """Returns device scalars ``(coef, norm)`` with ``coef = min(1, max_norm / (norm + 1e-6))``."""
    if N.use_native(sumsq):
        coef = torch.empty(1, dtype=torch.float32, device=sumsq.device)
        norm = torch.empty(1, dtype=torch.float32, device=sumsq.device)
        N.call("stb_clip_coef", sumsq.data_ptr(), float(max_norm or 0.0), coef.data_ptr(), norm.data_ptr(), N.stream_ptr())
        return coef, norm
    norm = sumsq.sqrt()
    coef = torch.clamp(max_norm / (norm + 1e-6), max=1.0) if max_norm and max_norm > 0 else torch.ones_like(norm)
    return coef, norm


def grad_clip_coef(g: torch.Tensor, max_norm: float, group=None) -> tuple[torch.Tensor, torch.Tensor]:
    """Global-norm clip coefficient of a flat gradient; ``group``: ranks whose squared norms add up
    (model-parallel groups -- fixes the reference's rank-local clipping)."""
    ss = sum_of_squares(g)
    if group is not None:
        import torch.distributed as dist

        if dist.get_world_size(group) > 1:
            dist.all_reduce(ss, group=group)
    return clip_coef_from_sumsq(ss, max_norm)


def fused_adamw_step(p, g, m, v, *, lr, betas, eps, weight_decay, step, clip_coef=None, grad_scale=1.0, hyper=None):
    """In-place AdamW on flat tensors: ``p`` bf16/fp32, ``g``/``m``/``v`` fp32.  ``hyper``: device tensor
    ``[lr, 1-β1^step, 1-β2^step]`` -- when given the launch takes those scalars from device memory (CUDA-graph safe)
    and ``lr`` / ``step`` are ignored."""
    if N.use_native(p):
        assert g.dtype == torch.float32 and m.dtype == torch.float32 and v.dtype == torch.float32
        if hyper is not None:
            N.call("stb_adamw_dyn", p.data_ptr(), int(p.dtype == torch.float32), g.data_ptr(), m.data_ptr(), v.data_ptr(),
                   p.numel(), float(betas[0]), float(betas[1]), float(eps), float(weight_decay), N.ptr(clip_coef),
                   float(grad_scale), hyper.data_ptr(), N.stream_ptr())
            return
        N.call("stb_adamw", p.data_ptr(), int(p.dtype == torch.float32), g.data_ptr(), m.data_ptr(), v.data_ptr(),
               p.numel(), float(lr), float(betas[0]), float(betas[1]), float(eps), float(weight_decay), int(step),
               N.ptr(clip_coef), float(grad_scale), N.stream_ptr())
        return
    gg = g.float() * grad_scale
    if clip_coef is not None:
        gg = gg * clip_coef
    m.mul_(betas[0]).add_(gg, alpha=1 - betas[0])
    v.mul_(betas[1]).addcmul_(gg, gg, value=1 - betas[1])
    bc1 = 1 - betas[0] ** step
    bc2 = 1 - betas[1] ** step
    denom = (v / bc2).sqrt_().add_(eps)
    pf = p.float() * (1 - lr * weight_decay) - lr * (m / bc1) / denom
    p.copy_(pf.to(p.dtype))
