"""Llama-2 architecture in plain ``torch.nn`` for the FSDP2 examples (reference: ``examples/FSDP2/llama2.py``): RMSNorm,
rotary embeddings (half-split layout, as in the HF checkpoints), grouped-query attention, SwiGLU MLP.  Named sizes follow
the published models; ``tiny`` / ``small`` exist so the examples run on CPU and on one GPU."""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class Llama2Config:
    vocab_size: int = 32000
    dim: int = 4096
    n_layers: int = 32
    n_heads: int = 32
    n_kv_heads: int = 32
    ffn_dim: int = 11008
    max_seq_len: int = 4096
    rope_theta: float = 10000.0
    norm_eps: float = 1e-5

    @staticmethod
    def named(name: str) -> "Llama2Config":
        table = {
            "tiny": dict(vocab_size=512, dim=128, n_layers=2, n_heads=4, n_kv_heads=2, ffn_dim=352, max_seq_len=256),
            "small": dict(vocab_size=32000, dim=768, n_layers=12, n_heads=12, n_kv_heads=12, ffn_dim=2048, max_seq_len=1024),
            "7b": dict(),
            "13b": dict(dim=5120, n_layers=40, n_heads=40, n_kv_heads=40, ffn_dim=13824),
            "70b": dict(dim=8192, n_layers=80, n_heads=64, n_kv_heads=8, ffn_dim=28672),
        }
        if name not in table:
            raise ValueError(f"unknown Llama-2 size {name}; choose from {sorted(table)}")
        return Llama2Config(**table[name])


class RMSNorm(nn.Module):
    def __init__(self, dim: int, eps: float):
        super().__init__()
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(dim))

    def forward(self, x):
        v = x.float()
        return (v * torch.rsqrt(v.pow(2).mean(-1, keepdim=True) + self.eps)).to(x.dtype) * self.weight


def rotary(seq: int, head_dim: int, theta: float, device=None):
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, device=device).float() / head_dim))
    ang = torch.outer(torch.arange(seq, device=device).float(), inv)
    ang = torch.cat([ang, ang], dim=-1)                    # half-split layout
    return ang.cos(), ang.sin()


def rotate(x, cos, sin):
    """x [B, H, S, D]"""
    d = x.shape[-1] // 2
    rot = torch.cat([-x[..., d:], x[..., :d]], dim=-1)
    S = x.shape[2]
    return x * cos[:S].to(x.dtype) + rot * sin[:S].to(x.dtype)


class Attention(nn.Module):
    def __init__(self, c: Llama2Config):
        super().__init__()
        self.hd = c.dim // c.n_heads
        self.q_proj = nn.Linear(c.dim, c.n_heads * self.hd, bias=False)
        self.k_proj = nn.Linear(c.dim, c.n_kv_heads * self.hd, bias=False)
        self.v_proj = nn.Linear(c.dim, c.n_kv_heads * self.hd, bias=False)
        self.o_proj = nn.Linear(c.n_heads * self.hd, c.dim, bias=False)

    def forward(self, x, cos, sin):
        B, S, _ = x.shape
        # head counts are inferred (-1): under tensor parallelism every rank sees its share of the heads
        q = self.q_proj(x).view(B, S, -1, self.hd).transpose(1, 2)
        k = self.k_proj(x).view(B, S, -1, self.hd).transpose(1, 2)
        v = self.v_proj(x).view(B, S, -1, self.hd).transpose(1, 2)
        q, k = rotate(q, cos, sin), rotate(k, cos, sin)
        y = F.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=q.shape[1] != k.shape[1])
        return self.o_proj(y.transpose(1, 2).reshape(B, S, -1))


class MLP(nn.Module):
    def __init__(self, c: Llama2Config):
        super().__init__()
        self.gate_proj = nn.Linear(c.dim, c.ffn_dim, bias=False)
        self.up_proj = nn.Linear(c.dim, c.ffn_dim, bias=False)
        self.down_proj = nn.Linear(c.ffn_dim, c.dim, bias=False)

    def forward(self, x):
        return self.down_proj(F.silu(self.gate_proj(x)) * self.up_proj(x))


class Block(nn.Module):
    def __init__(self, c: Llama2Config):
        super().__init__()
        self.input_layernorm = RMSNorm(c.dim, c.norm_eps)
        self.self_attn = Attention(c)
        self.post_attention_layernorm = RMSNorm(c.dim, c.norm_eps)
        self.mlp = MLP(c)

    def forward(self, x, cos, sin):
        x = x + self.self_attn(self.input_layernorm(x), cos, sin)
        return x + self.mlp(self.post_attention_layernorm(x))


class Llama2(nn.Module):
    def __init__(self, c: Llama2Config):
        super().__init__()
        self.config = c
        self.embed_tokens = nn.Embedding(c.vocab_size, c.dim)
        self.layers = nn.ModuleList([Block(c) for _ in range(c.n_layers)])
        self.norm = RMSNorm(c.dim, c.norm_eps)
        self.lm_head = nn.Linear(c.dim, c.vocab_size, bias=False)
        self._rope = None

    def init_weights(self, std: float = 0.02) -> None:
        """Called after ``to_empty`` when the model was built on the meta device."""
        for m in self.modules():
            if isinstance(m, (nn.Linear, nn.Embedding)):
                nn.init.normal_(m.weight, std=std)
            elif isinstance(m, RMSNorm):
                nn.init.ones_(m.weight)
        self._rope = None

    def forward(self, ids):
        if self._rope is None or self._rope[0].device != ids.device:
            c = self.config
            self._rope = rotary(c.max_seq_len, c.dim // c.n_heads, c.rope_theta, ids.device)
        cos, sin = self._rope
        h = self.embed_tokens(ids)
        for blk in self.layers:
            h = blk(h, cos, sin)
        return self.lm_head(self.norm(h))

    def num_params(self) -> int:
        return sum(p.numel() for p in self.parameters())
