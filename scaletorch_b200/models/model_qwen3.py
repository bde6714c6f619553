"""Qwen3 dense family (reference: ``scaletorch/models/model_qwen3.py:38-350``).

Explicit ``head_dim``, per-head q/k RMSNorm before RoPE (fused with RoPE and the attention node
here), attention projections initialised N(0, 0.02), optional tied embeddings, ``rope_theta`` 1e6
from the config.  Unlike the reference (``model_qwen3.py:260-269``) sequence parallelism works
for this family too.
"""
from __future__ import annotations

from .transformer import DecoderLM


class Qwen3(DecoderLM):
    """Qwen3 dense decoder: explicit ``head_dim``, per-head q / k RMSNorm before RoPE, optional tied embeddings."""
    model_type = "qwen3"
    qk_norm = True
    default_rope_theta = 1000000.0
    attn_init_std = 0.02

    def _head_dim(self, config) -> int:
        return getattr(config, "head_dim", None) or config.hidden_size // config.num_attention_heads
