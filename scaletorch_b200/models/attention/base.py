"""Shared machinery of the stand-alone attention modules (reference:
``scaletorch/models/attention/base.py:12-176``).

MHA, MQA and GQA are one computation with ``num_kv_heads = H, 1, G``; all of them run through
``ops.flash_attention`` (native GQA kernel, no K/V expansion) when no explicit mask / dropout / weight
output is requested, and through an eager fp32-softmax path otherwise.
"""
from __future__ import annotations

import math
from abc import ABC, abstractmethod

import torch
import torch.nn as nn

from ... import ops


def validate_attention_inputs(hidden_state: torch.Tensor, attention_mask: torch.Tensor | None, hidden_size: int) -> None:
    """Shape checks shared by the attention variants: ``hidden_state`` is ``[batch, seq, hidden_size]``, the mask has 2 to 4 dimensions."""
    if hidden_state.dim() != 3:
        raise ValueError(f"hidden_state must be [batch, seq, hidden], got {tuple(hidden_state.shape)}")
    if hidden_state.shape[-1] != hidden_size:
        raise ValueError(f"hidden_state last dim {hidden_state.shape[-1]} != hidden_size {hidden_size}")
    if attention_mask is not None and attention_mask.dim() not in (2, 3, 4):
        raise ValueError(f"attention_mask must have 2, 3 or 4 dims, got {attention_mask.dim()}")


class BaseAttention(nn.Module, ABC):
    """Common state and helpers of the stand-alone attention modules: head split / merge, mask broadcasting, and ``attend`` (flash kernel when possible, eager fp32 softmax when a mask, dropout or the weights are requested)."""
    def __init__(self, hidden_size: int, num_heads: int, dropout: float = 0.0, bias: bool = True) -> None:
        super().__init__()
        if hidden_size % num_heads != 0:
            raise ValueError(f"hidden_size ({hidden_size}) must be divisible by num_heads ({num_heads})")
        self.hidden_size, self.num_heads = hidden_size, num_heads
        self.head_dim = hidden_size // num_heads
        self.dropout_prob, self.bias = dropout, bias
        self.scale_factor = 1.0 / math.sqrt(self.head_dim)
        self.dropout = nn.Dropout(dropout)

    @abstractmethod
    def forward(self, hidden_state, attention_mask=None, return_attention_weights: bool = False):
        ...

    def split_head(self, x: torch.Tensor, num_heads: int | None = None) -> torch.Tensor:
        """``[B, S, h*d]`` → ``[B, S, h, d]`` (token-major, what the flash kernel consumes)."""
        B, S, _ = x.shape
        return x.view(B, S, num_heads or self.num_heads, self.head_dim)

    def combine_head(self, x: torch.Tensor) -> torch.Tensor:
        B, S = x.shape[:2]
        return x.reshape(B, S, -1)

    def apply_attention_mask(self, scores: torch.Tensor, attention_mask: torch.Tensor | None) -> torch.Tensor:
        """scores ``[B, H, Sq, Sk]``; mask is 1/True = keep, broadcast from 2-D ``[B, Sk]``, 3-D or 4-D."""
        if attention_mask is None:
            return scores
        m = attention_mask
        if m.dim() == 2:
            m = m[:, None, None, :]
        elif m.dim() == 3:
            m = m[:, None, :, :]
        return scores.masked_fill(m == 0, float("-inf"))

    def compute_attention_weights(self, q, k, attention_mask=None) -> torch.Tensor:
        """q ``[B,Sq,Hq,d]``, k ``[B,Sk,Hkv,d]`` → softmax weights ``[B,Hq,Sq,Sk]`` (fp32 softmax)."""
        g = q.shape[2] // k.shape[2]
        kf = k.repeat_interleave(g, dim=2) if g > 1 else k
        scores = torch.einsum("bqhd,bkhd->bhqk", q.float(), kf.float()) * self.scale_factor
        scores = self.apply_attention_mask(scores, attention_mask)
        return torch.softmax(scores, dim=-1)

    def attend(self, q, k, v, attention_mask=None, return_attention_weights: bool = False, causal: bool = False):
        """Core attention on token-major tensors; returns ``(out [B,Sq,Hq,d], weights | None)``."""
        fast = attention_mask is None and not return_attention_weights and (self.dropout_prob == 0.0 or not self.training)
        if fast:
            return ops.flash_attention(q, k, v, causal, self.scale_factor)[0], None
        w = self.compute_attention_weights(q, k, attention_mask)
        if causal:
            Sq, Sk = w.shape[-2:]
            tri = torch.ones(Sq, Sk, dtype=torch.bool, device=w.device).tril(Sk - Sq)
            w = torch.softmax(torch.log(w.clamp_min(1e-30)).masked_fill(~tri, float("-inf")), dim=-1)
        wd = self.dropout(w)
        g = q.shape[2] // v.shape[2]
        vf = v.repeat_interleave(g, dim=2) if g > 1 else v
        out = torch.einsum("bhqk,bkhd->bqhd", wd, vf.float()).to(q.dtype)
        return out, (w if return_attention_weights else None)

    def extra_repr(self) -> str:
        return f"hidden_size={self.hidden_size}, num_heads={self.num_heads}, head_dim={self.head_dim}, dropout={self.dropout_prob}"
