"""Multi-head attention (reference: ``scaletorch/models/attention/mha.py:9-141``)."""
from __future__ import annotations

import torch.nn as nn

from .base import BaseAttention, validate_attention_inputs


class MultiHeadAttention(BaseAttention):
    """Classic multi-head attention: one key / value head per query head."""
    def __init__(self, hidden_size: int, num_heads: int, dropout: float = 0.0, bias: bool = True, causal: bool = False):
        super().__init__(hidden_size, num_heads, dropout, bias)
        self.causal = causal
        self.qkv_proj = nn.Linear(hidden_size, 3 * hidden_size, bias=bias)   # fused q|k|v projection
        self.output_proj = nn.Linear(hidden_size, hidden_size, bias=bias)
        self._reset_parameters()

    def _reset_parameters(self) -> None:
        nn.init.xavier_uniform_(self.qkv_proj.weight)
        nn.init.xavier_uniform_(self.output_proj.weight)
        if self.bias:
            nn.init.zeros_(self.qkv_proj.bias)
            nn.init.zeros_(self.output_proj.bias)

    def forward(self, hidden_state, attention_mask=None, return_attention_weights: bool = False):
        validate_attention_inputs(hidden_state, attention_mask, self.hidden_size)
        B, S, _ = hidden_state.shape
        qkv = self.qkv_proj(hidden_state).view(B, S, 3, self.num_heads, self.head_dim)
        out, w = self.attend(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], attention_mask, return_attention_weights, self.causal)
        y = self.output_proj(self.combine_head(out))
        return (y, w) if return_attention_weights else y
