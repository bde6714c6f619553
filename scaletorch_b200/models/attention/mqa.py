"""Multi-query attention: all query heads share one K/V head (reference:
``scaletorch/models/attention/mqa.py:9-160``)."""
from __future__ import annotations

import torch.nn as nn

from .base import BaseAttention, validate_attention_inputs


class MultiQueryAttention(BaseAttention):
    """Multi-query attention: every query head attends over ONE shared key / value head."""
    def __init__(self, hidden_size: int, num_heads: int, dropout: float = 0.1, bias: bool = True, causal: bool = False):
        super().__init__(hidden_size, num_heads, dropout, bias)
        self.causal = causal
        self.q_proj = nn.Linear(hidden_size, hidden_size, bias=bias)
        self.kv_proj = nn.Linear(hidden_size, 2 * self.head_dim, bias=bias)  # one shared K head and one shared V head
        self.output_proj = nn.Linear(hidden_size, hidden_size, bias=bias)
        self._reset_parameters()

    def _reset_parameters(self) -> None:
        for m in (self.q_proj, self.kv_proj, self.output_proj):
            nn.init.xavier_uniform_(m.weight)
            if m.bias is not None:
                nn.init.zeros_(m.bias)

    def forward(self, hidden_state, attention_mask=None, return_attention_weights: bool = False):
        validate_attention_inputs(hidden_state, attention_mask, self.hidden_size)
        B, S, _ = hidden_state.shape
        q = self.split_head(self.q_proj(hidden_state))
        kv = self.kv_proj(hidden_state).view(B, S, 2, 1, self.head_dim)
        out, w = self.attend(q, kv[:, :, 0], kv[:, :, 1], attention_mask, return_attention_weights, self.causal)
        y = self.output_proj(self.combine_head(out))
        return (y, w) if return_attention_weights else y
