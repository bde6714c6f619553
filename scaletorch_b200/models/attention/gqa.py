"""Grouped-query attention (reference: ``scaletorch/models/attention/gqa.py:9-223``).

``num_kv_groups`` = number of K/V heads; each serves ``num_heads / num_kv_groups`` query heads.  The
reference expands K/V to the query head count before the matmul; here the kernel indexes the shared head.
"""
from __future__ import annotations

import torch.nn as nn

from .base import BaseAttention, validate_attention_inputs


class GroupQueryAttention(BaseAttention):
    """Grouped-query attention: ``num_heads`` query heads share ``num_kv_groups`` key / value heads (no K/V expansion on the kernel path)."""
    def __init__(self, hidden_size: int, num_heads: int, num_kv_groups: int, dropout: float = 0.1, bias: bool = True,
                 causal: bool = False):
        super().__init__(hidden_size, num_heads, dropout, bias)
        if num_heads % num_kv_groups != 0:
            raise ValueError(f"num_heads ({num_heads}) must be divisible by num_kv_groups ({num_kv_groups})")
        self.num_kv_groups = num_kv_groups
        self.heads_per_group = num_heads // num_kv_groups
        self.causal = causal
        self.q_proj = nn.Linear(hidden_size, hidden_size, bias=bias)
        self.kv_proj = nn.Linear(hidden_size, 2 * num_kv_groups * self.head_dim, bias=bias)
        self.output_proj = nn.Linear(hidden_size, hidden_size, bias=bias)
        self._reset_parameters()

    def _reset_parameters(self) -> None:
        for m in (self.q_proj, self.kv_proj, self.output_proj):
            nn.init.xavier_uniform_(m.weight)
            if m.bias is not None:
                nn.init.zeros_(m.bias)

    def split_head_grouped(self, x):
        B, S, _ = x.shape
        return x.view(B, S, self.num_kv_groups, self.head_dim)

    def forward(self, hidden_state, attention_mask=None, return_attention_weights: bool = False):
        validate_attention_inputs(hidden_state, attention_mask, self.hidden_size)
        B, S, _ = hidden_state.shape
        q = self.split_head(self.q_proj(hidden_state))
        kv = self.kv_proj(hidden_state).view(B, S, 2, self.num_kv_groups, self.head_dim)
        out, w = self.attend(q, kv[:, :, 0], kv[:, :, 1], attention_mask, return_attention_weights, self.causal)
        y = self.output_proj(self.combine_head(out))
        return (y, w) if return_attention_weights else y

    def extra_repr(self) -> str:
        return super().extra_repr() + f", num_kv_groups={self.num_kv_groups}"
