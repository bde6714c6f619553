"""Multi-head latent attention: low-rank query and shared low-rank K/V latents (reference:
``scaletorch/models/attention/mla.py:9-152`` -- like the reference, without RoPE / a decoupled rope head)."""
from __future__ import annotations

import torch.nn as nn

from .base import BaseAttention, validate_attention_inputs


class MultiHeadLatentAttention(BaseAttention):
    """Multi-head latent attention: queries through a ``q_latent_size`` bottleneck, keys and values expanded from one shared ``kv_latent_size`` latent (the tensor an inference cache would keep)."""
    def __init__(self, hidden_size: int, num_heads: int, q_latent_size: int, kv_latent_size: int, dropout: float = 0.0,
                 bias: bool = True, causal: bool = False):
        super().__init__(hidden_size, num_heads, dropout, bias)
        self.q_latent_size, self.kv_latent_size, self.causal = q_latent_size, kv_latent_size, causal
        self.q_down_proj = nn.Linear(hidden_size, q_latent_size, bias=bias)
        self.q_up_proj = nn.Linear(q_latent_size, hidden_size, bias=bias)
        self.kv_down_proj = nn.Linear(hidden_size, kv_latent_size, bias=bias)
        self.kv_up_proj = nn.Linear(kv_latent_size, 2 * hidden_size, bias=bias)   # fused k_up | v_up
        self.output_proj = nn.Linear(hidden_size, hidden_size, bias=bias)

    def compress_kv(self, hidden_state):
        """The latent that an inference cache would store (``[B, S, kv_latent_size]``)."""
        return self.kv_down_proj(hidden_state)

    def forward(self, hidden_state, attention_mask=None, return_attention_weights: bool = False):
        validate_attention_inputs(hidden_state, attention_mask, self.hidden_size)
        B, S, _ = hidden_state.shape
        q = self.split_head(self.q_up_proj(self.q_down_proj(hidden_state)))
        kv = self.kv_up_proj(self.compress_kv(hidden_state)).view(B, S, 2, self.num_heads, self.head_dim)
        out, w = self.attend(q, kv[:, :, 0], kv[:, :, 1], attention_mask, return_attention_weights, self.causal)
        y = self.output_proj(self.combine_head(out))
        return (y, w) if return_attention_weights else y
