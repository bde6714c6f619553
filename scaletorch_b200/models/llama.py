"""Llama family (reference: ``scaletorch/models/llama.py:60-556``).

``head_dim = hidden/heads``, no QK-norm, ``rope_theta`` default 500000, uniform ±1/√fan_in init.
Sequence parallelism is a constructor-time property of the TP layers (the reference brackets the
layer with gather/scatter calls that double-reduce).
"""
from __future__ import annotations

from .transformer import DecoderLM


class Llama(DecoderLM):
    """Llama-family decoder: no QK-norm, ``rope_theta`` 5e5 by default, uniform fan-in initialisation."""
    model_type = "llama"
    qk_norm = False
    default_rope_theta = 500000.0
    attn_init_std = None  # uniform(±1/sqrt(fan_in))

    def _head_dim(self, config) -> int:
        return getattr(config, "head_dim", None) or config.hidden_size // config.num_attention_heads
