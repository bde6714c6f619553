"""Shared decoder-only transformer used by Llama, Qwen3 and Qwen3-MoE.

Reference model code: ``scaletorch/models/llama.py:60-556``, ``model_qwen3.py:38-350``,
``model_qwen3_moe.py:295-409``.  What is kept: module names (``embedding``, ``decoder_layers``,
``final_norm``, ``final_proj``; per layer ``input_layernorm``, ``attention``, ``post_attention_layernorm``,
``mlp`` / ``moe``), ``reset_parameters()``, ``forward(input_ids, attention_mask, position_ids,
gradient_checkpointing)`` returning full logits, tied embeddings by parameter sharing.

What is designed differently for B200:
* q/k/v and gate/up are single fused projections (one GEMM each, two TP collectives per layer in
  backward instead of five);
* activations are token-major ``[T, H]``; the residual stream is carried as ``(h, delta)`` so every
  residual add is fused into the following RMSNorm kernel;
* QK-norm + RoPE + attention (+ their backward) are one autograd node writing a packed ``dqkv``;
* the LM head is never materialised in fp32: ``loss()`` runs the fused vocab-parallel linear+CE;
* parallelism comes from an explicit :class:`ParallelContext` (TP/SP layers, ring attention under
  CP, expert parallel MoE), not from environment variables read in constructors.
"""
from __future__ import annotations

import math
from typing import Any

import torch
import torch.nn as nn
from torch.utils.checkpoint import checkpoint as torch_checkpoint

from .. import ops
from ..parallel.context import ParallelContext
from ..parallel.tensor_parallel import tp_comms
from ..parallel.tensor_parallel.tensor_parallel import (ColumnParallelLinear, RowParallelLinear,
                                                        VocabParallelEmbedding)
from .attention_utils import RMSNorm, get_attention_backend, resolve_attention_backend_name


class Attention(nn.Module):
    """Self-attention of the decoder: ONE fused q|k|v projection (column-parallel), optional per-head QK-norm fused with RoPE, flash / ring attention with native GQA, row-parallel output projection."""
    def __init__(self, config: Any, layer_idx: int, ctx: ParallelContext, *, qk_norm: bool, head_dim: int,
                 init_std: float | None):
        super().__init__()
        self.layer_idx = layer_idx
        self.ctx = ctx
        self.hidden_size = config.hidden_size
        self.num_heads = config.num_attention_heads
        self.num_key_values = config.num_key_value_heads
        self.head_dim = head_dim
        tp = ctx.tp_size
        if self.num_heads % tp != 0:
            raise ValueError(f"num_attention_heads ({self.num_heads}) must be divisible by tp_world_size ({tp})")
        if self.num_key_values % tp != 0:
            raise ValueError(f"num_key_value_heads ({self.num_key_values}) must be divisible by tp_world_size ({tp})")
        if self.num_heads % self.num_key_values != 0:
            raise ValueError("num_attention_heads must be a multiple of num_key_value_heads")
        self.num_local_heads = self.num_heads // tp
        self.num_local_kv_heads = self.num_key_values // tp
        self.eps = getattr(config, "rms_norm_eps", 1e-6)
        kw = dict(group=ctx.tp_group, tp_size=ctx.tp_size, tp_rank=ctx.tp_rank, sequence_parallel=ctx.sp,
                  fused_comm=ctx.fused_comm, init_std=init_std)
        self.qkv_proj = ColumnParallelLinear(self.hidden_size, (self.num_heads + 2 * self.num_key_values) * head_dim,
                                             bias=getattr(config, "attention_bias", False), **kw)
        self.out_proj = RowParallelLinear(self.num_heads * head_dim, self.hidden_size, **kw)
        self.q_norm = RMSNorm(head_dim, self.eps) if qk_norm else None
        self.k_norm = RMSNorm(head_dim, self.eps) if qk_norm else None
        self._attn_backend_name = resolve_attention_backend_name(ctx.cp_size > 1, ctx.use_flash_attention)

    def reset_parameters(self) -> None:
        self.qkv_proj.reset_parameters()
        self.out_proj.reset_parameters()
        if self.q_norm is not None:
            self.q_norm.reset_parameters()
            self.k_norm.reset_parameters()

    def forward(self, x: torch.Tensor, B: int, S: int, cos: torch.Tensor, sin: torch.Tensor, pos_offset: int = 0,
                positions: torch.Tensor | None = None):
        """x: ``[T(/tp), H]`` → ``[T(/tp), H]`` with ``T = B*S`` local (CP-sharded) tokens.  ``positions [T]``: explicit
        token positions (zig-zag context parallelism), otherwise ``pos_offset + t mod S``."""
        qkv = self.qkv_proj(x)  # [T, (Hq_l + 2 Hkv_l) * D]
        attn_fn = None
        if self._attn_backend_name != "flash":
            attn_fn = get_attention_backend(self._attn_backend_name)
        a = ops.qkv_attention(
            qkv, self.q_norm.weight if self.q_norm is not None else None,
            self.k_norm.weight if self.k_norm is not None else None, cos, sin, B, S,
            self.num_local_heads, self.num_local_kv_heads, self.head_dim, pos_offset, self.eps,
            causal=True, scale=1.0 / math.sqrt(self.head_dim), attn_fn=attn_fn, positions=positions)
        return self.out_proj(a.reshape(B * S, self.num_local_heads * self.head_dim))


class MLP(nn.Module):
    """SwiGLU feed-forward with a fused ``gate|up`` projection."""

    def __init__(self, config: Any, ctx: ParallelContext, init_std: float | None = None):
        super().__init__()
        self.hidden_size = config.hidden_size
        self.intermediate_size = config.intermediate_size
        kw = dict(group=ctx.tp_group, tp_size=ctx.tp_size, tp_rank=ctx.tp_rank, sequence_parallel=ctx.sp,
                  fused_comm=ctx.fused_comm, init_std=init_std)
        bias = getattr(config, "mlp_bias", False)
        self.gate_up_proj = ColumnParallelLinear(self.hidden_size, 2 * self.intermediate_size, bias=bias, **kw)
        self.down_proj = RowParallelLinear(self.intermediate_size, self.hidden_size, bias=bias, **kw)

    def reset_parameters(self) -> None:
        self.gate_up_proj.reset_parameters()
        self.down_proj.reset_parameters()

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.down_proj(ops.swiglu(self.gate_up_proj(x)))


class DecoderLayer(nn.Module):
    """Pre-norm decoder layer on the ``(h, delta)`` residual stream: the add of the previous branch output is fused into this layer's RMSNorm; the feed-forward is the SwiGLU MLP or the module passed as ``ffn`` (MoE)."""
    def __init__(self, config: Any, layer_idx: int, ctx: ParallelContext, *, qk_norm: bool, head_dim: int,
                 init_std: float | None, ffn: nn.Module | None = None):
        super().__init__()
        self.layer_idx = layer_idx
        eps = getattr(config, "rms_norm_eps", 1e-6)
        self.input_layernorm = RMSNorm(config.hidden_size, eps)
        self.post_attention_layernorm = RMSNorm(config.hidden_size, eps)
        self.attention = Attention(config, layer_idx, ctx, qk_norm=qk_norm, head_dim=head_dim, init_std=init_std)
        if ffn is None:
            self.mlp = MLP(config, ctx)
            self._ffn_name = "mlp"
        else:
            self.moe = ffn
            self._ffn_name = "moe"

    @property
    def ffn(self) -> nn.Module:
        return getattr(self, self._ffn_name)

    def reset_parameters(self) -> None:
        self.input_layernorm.reset_parameters()
        self.post_attention_layernorm.reset_parameters()
        self.attention.reset_parameters()
        self.ffn.reset_parameters()

    def forward(self, h: torch.Tensor, delta: torch.Tensor | None, B: int, S: int, cos, sin, pos_offset: int = 0,
                positions: torch.Tensor | None = None):
        """``(h, delta)`` is the residual stream and the not-yet-added branch output."""
        if delta is None:
            y = ops.rmsnorm(h, self.input_layernorm.weight, self.input_layernorm.eps)
        else:
            h, y = ops.add_rmsnorm(delta, h, self.input_layernorm.weight, self.input_layernorm.eps)
        a = self.attention(y, B, S, cos, sin, pos_offset, positions)
        h, y = ops.add_rmsnorm(a, h, self.post_attention_layernorm.weight, self.post_attention_layernorm.eps)
        return h, self.ffn(y)


class FinalProjection(ColumnParallelLinear):
    """Vocabulary projection, vocab-sharded under TP (``gather_output`` restores full logits)."""


class DecoderLM(nn.Module):
    """Base class; subclasses pick ``qk_norm`` / ``head_dim`` / init / FFN type."""

    model_type = "decoder"
    qk_norm = False
    default_rope_theta = 500000.0
    attn_init_std: float | None = None

    def __init__(self, config: Any, ctx: ParallelContext | None = None) -> None:
        super().__init__()
        self.ctx = ctx if ctx is not None else ParallelContext.from_pgm()
        ctx = self.ctx
        self.config = self.model_config = config
        self.vocab_size = config.vocab_size
        self.hidden_size = config.hidden_size
        self.num_heads = config.num_attention_heads
        self.num_key_values = config.num_key_value_heads
        self.head_dim = self._head_dim(config)
        self.max_position_embeddings = config.max_position_embeddings
        self.num_layers = config.num_hidden_layers
        self.tie_word_embeddings = bool(getattr(config, "tie_word_embeddings", False))
        self._stb_tp_constructed = True

        tpkw = dict(group=ctx.tp_group, tp_size=ctx.tp_size, tp_rank=ctx.tp_rank)
        self.embedding = VocabParallelEmbedding(self.vocab_size, self.hidden_size, sequence_parallel=ctx.sp, **tpkw)
        self.decoder_layers = nn.ModuleList([self._make_layer(config, i) for i in range(self.num_layers)])
        self.final_norm = RMSNorm(self.hidden_size, getattr(config, "rms_norm_eps", 1e-6))
        self.final_proj = FinalProjection(self.hidden_size, self.vocab_size, bias=False, gather_output=True,
                                          sequence_parallel=ctx.sp, fused_comm=ctx.fused_comm, **tpkw)
        if self.tie_word_embeddings:
            self.final_proj.weight = self.embedding.weight  # same shard shape [V/tp, H]
            self.embedding.weight._stb_expected_grads = 2   # LM-head wgrad + embedding scatter-add

        rope_theta = getattr(config, "rope_theta", None)
        if rope_theta is None:
            rp = getattr(config, "rope_parameters", None) or {}
            rope_theta = rp.get("rope_theta", self.default_rope_theta) if isinstance(rp, dict) else self.default_rope_theta
        cos, sin = ops.get_cos_sin(config.max_position_embeddings, self.head_dim, float(rope_theta), device="cpu")
        self.register_buffer("cos", cos, persistent=False)
        self.register_buffer("sin", sin, persistent=False)

    # ------------------------------------------------------------------ construction hooks
    def _head_dim(self, config) -> int:
        return config.hidden_size // config.num_attention_heads

    def _make_ffn(self, config, layer_idx: int) -> nn.Module | None:
        return None  # dense SwiGLU MLP

    def _make_layer(self, config, layer_idx: int) -> DecoderLayer:
        return DecoderLayer(config, layer_idx, self.ctx, qk_norm=self.qk_norm, head_dim=self.head_dim,
                            init_std=self.attn_init_std, ffn=self._make_ffn(config, layer_idx))

    def reset_parameters(self) -> None:
        self.embedding.reset_parameters()
        for layer in self.decoder_layers:
            layer.reset_parameters()
        self.final_norm.reset_parameters()
        if not self.tie_word_embeddings:
            self.final_proj.reset_parameters()

    def set_sequence_parallel(self, enabled: bool) -> None:
        enabled = enabled and self.ctx.tp_size > 1
        self.ctx.sequence_parallel = enabled
        for m in self.modules():
            if isinstance(m, (ColumnParallelLinear, RowParallelLinear, VocabParallelEmbedding)):
                m.sequence_parallel = enabled

    def _apply(self, fn, *args, **kwargs):
        # keep the RoPE tables in fp32 when the module is cast to bf16
        cos, sin = self.cos, self.sin
        out = super()._apply(fn, *args, **kwargs)
        if cos.dtype == torch.float32 and self.cos.dtype != torch.float32:
            self.cos = cos.to(self.cos.device)
            self.sin = sin.to(self.sin.device)
        return out

    # ------------------------------------------------------------------ forward pieces
    def embed(self, input_ids: torch.Tensor) -> torch.Tensor:
        """``[B, S]`` ids → token-major ``[T(/tp), H]``."""
        x = self.embedding(input_ids)
        return x.reshape(-1, self.hidden_size)

    def run_layers(self, h, delta, B: int, S: int, layers=None, gradient_checkpointing: bool = False):
        pos_offset = self.ctx.cp_rank * S if self.ctx.cp_size > 1 else 0
        positions = None
        if self.ctx.cp_size > 1:
            from ..parallel.context_parallel import zigzag

            if zigzag.enabled():   # this rank holds chunks cp_rank and 2·cp−1−cp_rank: RoPE needs explicit positions
                positions = zigzag.local_positions(S, self.ctx.cp_size, self.ctx.cp_rank, h.device).repeat(B)
        layers = self.decoder_layers if layers is None else layers
        for layer in layers:
            if gradient_checkpointing and torch.is_grad_enabled():
                h, delta = torch_checkpoint(layer, h, delta, B, S, self.cos, self.sin, pos_offset, positions,
                                            use_reentrant=False)
            else:
                h, delta = layer(h, delta, B, S, self.cos, self.sin, pos_offset, positions)
        return h, delta

    def finalize(self, h, delta):
        """Adds the pending branch output and applies the final norm → ``[T(/tp), H]``."""
        if delta is None:
            return ops.rmsnorm(h, self.final_norm.weight, self.final_norm.eps)
        return ops.add_rmsnorm(delta, h, self.final_norm.weight, self.final_norm.eps)[1]

    def hidden_states(self, input_ids, gradient_checkpointing: bool = False):
        B, S = input_ids.shape
        h, delta = self.run_layers(self.embed(input_ids), None, B, S, gradient_checkpointing=gradient_checkpointing)
        return self.finalize(h, delta)

    def forward(self, input_ids, attention_mask=None, position_ids=None, gradient_checkpointing: bool = False):
        """Full logits ``[B, S, V]`` (reference-compatible; prefer :meth:`loss` for training)."""
        B, S = input_ids.shape
        y = self.hidden_states(input_ids, gradient_checkpointing)
        logits = self.final_proj(y)  # column-parallel + gather (all-gathers tokens first under SP)
        return logits.view(B, S, -1)

    def loss(self, input_ids, target_ids, gradient_checkpointing: bool = False, loss_scale: float = 1.0,
             ignore_index: int = -100) -> torch.Tensor:
        """Mean token cross-entropy via the fused (vocab-parallel) linear+CE -- no fp32 / full-vocab logits."""
        y = self.hidden_states(input_ids, gradient_checkpointing)
        return self.lm_loss(y, target_ids, loss_scale, ignore_index)

    def lm_loss(self, y, target_ids, loss_scale: float = 1.0, ignore_index: int = -100) -> torch.Tensor:
        ctx = self.ctx
        if ctx.sp:
            y = tp_comms.gather_from_sequence_parallel_region(y, ctx.tp_group)
        elif ctx.tp_size > 1:
            y = tp_comms.copy_to_tensor_parallel_region(y, ctx.tp_group)
        vocab_start = ctx.tp_rank * (self.vocab_size // ctx.tp_size)
        return ops.linear_cross_entropy(y, self.final_proj.weight, target_ids.reshape(-1), loss_scale=loss_scale,
                                        ignore_index=ignore_index,
                                        tp_group=ctx.tp_group if ctx.tp_size > 1 else None, vocab_start=vocab_start)

    def get_aux_loss(self) -> torch.Tensor | None:
        return None

    @torch.no_grad()
    def generate(self, input_ids: torch.Tensor, max_new_tokens: int, temperature: float = 0.0, top_k: int | None = None,
                 eos_token_id: int | None = None, generator: torch.Generator | None = None) -> torch.Tensor:
        """Autoregressive continuation of ``input_ids [B, S]`` (sanity checks, evaluation samples, parity tests against
        ``transformers.generate``).  ``temperature == 0`` is greedy.  No KV cache: every step re-runs the prompt through
        the training kernels, which is the point -- it exercises exactly the forward the trainer uses; the context is
        clipped to ``max_position_embeddings``.  Unsharded or tensor-parallel models (every TP rank returns the same
        tokens); not for pipeline stages."""
        was_training = self.training
        self.eval()
        max_ctx = int(getattr(self.config, "max_position_embeddings", 1 << 30))
        done = torch.zeros(input_ids.shape[0], dtype=torch.bool, device=input_ids.device)
        out = input_ids
        for _ in range(max_new_tokens):
            logits = self(out[:, -max_ctx:])[:, -1, :].float()
            if temperature and temperature > 0:
                logits = logits / temperature
                if top_k is not None:
                    kth = torch.topk(logits, min(top_k, logits.shape[-1]))[0][:, -1:]
                    logits = logits.masked_fill(logits < kth, float("-inf"))
                nxt = torch.multinomial(torch.softmax(logits, dim=-1), 1, generator=generator)
            else:
                nxt = logits.argmax(dim=-1, keepdim=True)
            if eos_token_id is not None:
                nxt = torch.where(done.unsqueeze(1), torch.full_like(nxt, eos_token_id), nxt)
                done |= nxt.squeeze(1) == eos_token_id
            out = torch.cat([out, nxt], dim=1)
            if eos_token_id is not None and bool(done.all()):
                break
        self.train(was_training)
        return out
