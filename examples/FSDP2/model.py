"""A small Llama-style decoder used by the FSDP2 examples (plain ``torch.nn`` on purpose: these examples demonstrate the
*upstream* sharding APIs next to the framework's own 5-D parallel engine; reference: ``examples/FSDP2/{model,llama2}.py``)."""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class ModelArgs:
    vocab_size: int = 4096
    dim: int = 256
    n_layers: int = 4
    n_heads: int = 8
    n_kv_heads: int = 4
    ffn_dim: int = 704
    max_seq_len: int = 512
    rope_theta: float = 10000.0
    norm_eps: float = 1e-5


def rope_tables(seq: int, head_dim: int, theta: float, device=None):
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, device=device).float() / head_dim))
    ang = torch.outer(torch.arange(seq, device=device).float(), inv)
    return torch.cos(ang), torch.sin(ang)


def apply_rope(x, cos, sin):
    """x [B, H, S, D]; interleaved-pair rotation."""
    x1, x2 = x[..., 0::2], x[..., 1::2]
    cos, sin = cos[: x.shape[2]].to(x.dtype), sin[: x.shape[2]].to(x.dtype)
    return torch.stack([x1 * cos - x2 * sin, x1 * sin + x2 * cos], dim=-1).flatten(-2)


class Attention(nn.Module):
    def __init__(self, a: ModelArgs):
        super().__init__()
        self.n_heads, self.n_kv, self.hd = a.n_heads, a.n_kv_heads, a.dim // a.n_heads
        self.wq = nn.Linear(a.dim, a.n_heads * self.hd, bias=False)
        self.wk = nn.Linear(a.dim, a.n_kv_heads * self.hd, bias=False)
        self.wv = nn.Linear(a.dim, a.n_kv_heads * self.hd, bias=False)
        self.wo = nn.Linear(a.n_heads * self.hd, a.dim, bias=False)

    def forward(self, x, cos, sin):
        B, S, _ = x.shape
        q = self.wq(x).view(B, S, -1, self.hd).transpose(1, 2)       # -1: head count is local under tensor parallelism
        k = self.wk(x).view(B, S, -1, self.hd).transpose(1, 2)
        v = self.wv(x).view(B, S, -1, self.hd).transpose(1, 2)
        q, k = apply_rope(q, cos, sin), apply_rope(k, cos, sin)
        y = F.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=q.shape[1] != k.shape[1])
        return self.wo(y.transpose(1, 2).reshape(B, S, -1))


class FeedForward(nn.Module):
    def __init__(self, a: ModelArgs):
        super().__init__()
        self.w1 = nn.Linear(a.dim, a.ffn_dim, bias=False)
        self.w3 = nn.Linear(a.dim, a.ffn_dim, bias=False)
        self.w2 = nn.Linear(a.ffn_dim, a.dim, bias=False)

    def forward(self, x):
        return self.w2(F.silu(self.w1(x)) * self.w3(x))


class Block(nn.Module):
    def __init__(self, a: ModelArgs):
        super().__init__()
        self.attention_norm = nn.RMSNorm(a.dim, eps=a.norm_eps)
        self.attention = Attention(a)
        self.ffn_norm = nn.RMSNorm(a.dim, eps=a.norm_eps)
        self.feed_forward = FeedForward(a)

    def forward(self, x, cos, sin):
        x = x + self.attention(self.attention_norm(x), cos, sin)
        return x + self.feed_forward(self.ffn_norm(x))


class Transformer(nn.Module):
    def __init__(self, a: ModelArgs):
        super().__init__()
        self.args = a
        self.tok_embeddings = nn.Embedding(a.vocab_size, a.dim)
        self.layers = nn.ModuleList(Block(a) for _ in range(a.n_layers))
        self.norm = nn.RMSNorm(a.dim, eps=a.norm_eps)
        self.output = nn.Linear(a.dim, a.vocab_size, bias=False)

    def init_weights(self):
        for m in self.modules():
            if isinstance(m, (nn.Linear, nn.Embedding)):
                nn.init.normal_(m.weight, std=0.02)
            elif isinstance(m, nn.RMSNorm):
                nn.init.ones_(m.weight)

    def forward(self, tokens):
        cos, sin = rope_tables(tokens.shape[1], self.args.dim // self.args.n_heads, self.args.rope_theta, tokens.device)
        h = self.tok_embeddings(tokens)
        for layer in self.layers:
            h = layer(h, cos, sin)
        return self.output(self.norm(h))
