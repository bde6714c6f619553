"""nanoGPT-style GPT with capacity-based Mixture-of-Experts layers.

Reference: ``scaletorch/models/moe.py:41-903`` (``GPTConfig``, noisy top-k ``Router`` with capacity
``floor(k · cf · T / E)``, auxiliary + z losses, ``MLPExperts``, ``MOELayer``, ``GPT`` with generate / crop /
``estimate_mfu``, ``analyze_moe_usage``, ``get_moe_layer_info``).

Design difference: the reference builds a dense ``[T, E, C]`` dispatch tensor and einsums through it; here
tokens that fit their expert's capacity are *sorted by expert* and run through the same ragged grouped GEMM
as Qwen3-MoE (``ops.grouped``), then scattered back with their gate weights -- O(T·k·H) memory instead of
O(T·E·C), same drop semantics (position-in-expert ≥ capacity ⇒ dropped, earlier top-k choices first).
Attention goes through ``ops.flash_attention``.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Any

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops
from ..ops.grouped import grouped_linear


@dataclass
class GPTConfig:
    """nanoGPT-style hyper-parameters plus the MoE switches (which layers are sparse, experts, top-k, capacity, auxiliary / z losses, noisy routing); validated in ``__post_init__``."""
    block_size: int = 1024
    vocab_size: int = 50304
    n_layer: int = 12
    n_head: int = 12
    n_embd: int = 768
    dropout: float = 0.0
    bias: bool = True
    use_moe: bool = False
    moe_layers: list[int] | None = None
    n_experts: int = 8
    top_k: int = 2
    capacity_factor: float = 1.25
    use_aux_loss: bool = False
    use_router_z_loss: bool = False
    use_noisy_top_k: bool = True
    aux_loss_weight: float = 0.01
    router_z_loss_weight: float = 0.01
    train_capacity: float = 1.25
    eval_capacity: float = 1.25
    min_capacity: float = 0.0
    use_switch_tfm_init: bool = False
    switch_tfm_init_scale: float = 1.0
    use_optimized_routing: bool = True
    use_einsum_aggregation: bool = True
    expert_capacity_roundup: bool = True

    def __post_init__(self) -> None:
        for name in ("block_size", "vocab_size", "n_layer", "n_head", "n_embd"):
            if getattr(self, name) <= 0:
                raise ValueError(f"{name} must be positive, got {getattr(self, name)}")
        if self.n_embd % self.n_head != 0:
            raise ValueError(f"n_embd ({self.n_embd}) must be divisible by n_head ({self.n_head})")
        if not 0.0 <= self.dropout <= 1.0:
            raise ValueError(f"dropout must be in [0, 1], got {self.dropout}")
        if self.use_moe:
            if self.n_experts <= 0:
                raise ValueError(f"n_experts must be positive, got {self.n_experts}")
            if not 1 <= self.top_k <= self.n_experts:
                raise ValueError(f"top_k ({self.top_k}) must be between 1 and n_experts ({self.n_experts})")
            if self.capacity_factor <= 0:
                raise ValueError(f"capacity_factor must be positive, got {self.capacity_factor}")
        if self.switch_tfm_init_scale <= 0:
            raise ValueError(f"switch_tfm_init_scale must be positive, got {self.switch_tfm_init_scale}")

    def is_moe_layer(self, idx: int) -> bool:
        if not self.use_moe:
            return False
        if self.moe_layers is not None:
            return idx in self.moe_layers
        return idx % 2 == 1  # default pattern: every other layer


class LayerNorm(nn.Module):
    """LayerNorm with an optional bias (``F.layer_norm`` has no bias switch)."""
    def __init__(self, ndim: int, bias: bool) -> None:
        super().__init__()
        self.weight = nn.Parameter(torch.ones(ndim))
        self.bias = nn.Parameter(torch.zeros(ndim)) if bias else None

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return F.layer_norm(x, self.weight.shape, self.weight, self.bias, 1e-5)


class CausalSelfAttention(nn.Module):
    """Fused-qkv causal self-attention through ``ops.flash_attention``."""
    def __init__(self, config: GPTConfig) -> None:
        super().__init__()
        self.n_head, self.n_embd, self.dropout = config.n_head, config.n_embd, config.dropout
        self.c_attn = nn.Linear(config.n_embd, 3 * config.n_embd, bias=config.bias)
        self.c_proj = nn.Linear(config.n_embd, config.n_embd, bias=config.bias)
        self.resid_dropout = nn.Dropout(config.dropout)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        B, T, C = x.shape
        qkv = self.c_attn(x).view(B, T, 3, self.n_head, C // self.n_head)
        y = ops.flash_attention(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], True)[0]
        return self.resid_dropout(self.c_proj(y.reshape(B, T, C)))


class GELUMLP(nn.Module):
    """The dense feed-forward block: 4x expansion, GELU, projection, dropout."""
    def __init__(self, config: GPTConfig) -> None:
        super().__init__()
        self.c_fc = nn.Linear(config.n_embd, 4 * config.n_embd, bias=config.bias)
        self.c_proj = nn.Linear(4 * config.n_embd, config.n_embd, bias=config.bias)
        self.dropout = nn.Dropout(config.dropout)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.dropout(self.c_proj(F.gelu(self.c_fc(x))))


class Block(nn.Module):
    """Pre-norm transformer block with the dense MLP."""
    def __init__(self, config: GPTConfig) -> None:
        super().__init__()
        self.ln_1, self.ln_2 = LayerNorm(config.n_embd, config.bias), LayerNorm(config.n_embd, config.bias)
        self.attn, self.mlp = CausalSelfAttention(config), GELUMLP(config)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = x + self.attn(self.ln_1(x))
        return x + self.mlp(self.ln_2(x))


class MLPExperts(nn.Module):
    """``n_experts`` GELU MLPs as stacked weights ``c_fc [E, 4H, H]`` / ``c_proj [E, H, 4H]`` (+ biases)."""

    def __init__(self, config: GPTConfig) -> None:
        super().__init__()
        E, H = config.n_experts, config.n_embd
        self.config = config
        self.c_fc = nn.Parameter(torch.empty(E, 4 * H, H))
        self.c_proj = nn.Parameter(torch.empty(E, H, 4 * H))
        self.fc_bias = nn.Parameter(torch.zeros(E, 4 * H)) if config.bias else None
        self.proj_bias = nn.Parameter(torch.zeros(E, H)) if config.bias else None
        self.dropout = nn.Dropout(config.dropout)
        self._init_weights()

    def _init_weights(self) -> None:
        cfg = self.config
        if cfg.use_switch_tfm_init:
            for w in (self.c_fc, self.c_proj):
                std = math.sqrt(cfg.switch_tfm_init_scale / w.shape[-1])
                nn.init.trunc_normal_(w, mean=0.0, std=std, a=-2 * std, b=2 * std)
        else:
            nn.init.normal_(self.c_fc, std=0.02)
            nn.init.normal_(self.c_proj, std=0.02 / math.sqrt(2 * cfg.n_layer))

    def forward(self, x_sorted: torch.Tensor, tokens_per_expert: torch.Tensor) -> torch.Tensor:
        """x_sorted ``[R, H]`` grouped by expert → ``[R, H]``."""
        h = grouped_linear(x_sorted, self.c_fc, tokens_per_expert)
        if self.fc_bias is not None:
            # expert id of every row; rows past sum(tokens_per_expert) are dropped tokens (their output is
            # discarded by the caller), give them the last expert's bias
            bounds = torch.cumsum(tokens_per_expert, 0)
            rows = torch.arange(x_sorted.shape[0], device=x_sorted.device)
            eid = torch.bucketize(rows, bounds, right=True).clamp_(max=self.c_fc.shape[0] - 1)
            h = h + self.fc_bias[eid]
        h = F.gelu(h)
        y = grouped_linear(h, self.c_proj, tokens_per_expert)
        if self.proj_bias is not None:
            y = y + self.proj_bias[eid]
        return self.dropout(y)


class Router(nn.Module):
    """Noisy top-k router with a per-expert capacity ``floor(k * cf * T / E)``: returns gate weights, expert ids, the keep mask of pairs that fit their expert's queue (earlier choices first), the capacity and the auxiliary / z loss."""
    def __init__(self, config: GPTConfig) -> None:
        super().__init__()
        self.config = config
        self.n_experts, self.top_k = config.n_experts, config.top_k
        self.w_g = nn.Linear(config.n_embd, config.n_experts, bias=False)
        self.w_noise = nn.Linear(config.n_embd, config.n_experts, bias=False) if config.use_noisy_top_k else None

    def _compute_gate_scores(self, x: torch.Tensor) -> torch.Tensor:
        logits = self.w_g(x).float()
        if self.w_noise is not None and self.training:
            noise = F.softplus(self.w_noise(x).float())
            logits = logits + noise * torch.randn_like(noise)
        return logits

    def _select_experts(self, logits: torch.Tensor):
        top_logits, top_idx = torch.topk(logits, self.top_k, dim=-1)
        return F.softmax(top_logits, dim=-1), top_idx

    def _compute_expert_capacity(self, tokens_per_batch: int) -> int:
        cf = self.config.train_capacity if self.training else self.config.eval_capacity
        cap = math.floor(self.top_k * cf * tokens_per_batch / self.n_experts)
        if self.config.expert_capacity_roundup:
            cap += cap % 2
        return max(int(cap), int(self.config.min_capacity), 1)

    def compute_router_z_loss(self, logits: torch.Tensor) -> torch.Tensor:
        return torch.logsumexp(logits, dim=-1).pow(2).mean()

    def compute_aux_loss(self, probs: torch.Tensor, top_idx: torch.Tensor) -> torch.Tensor:
        """Switch-style load balancing: ``E · Σ_e f_e · p_e``."""
        T = probs.shape[0]
        f = torch.bincount(top_idx.reshape(-1), minlength=self.n_experts).to(probs.dtype) / (T * self.top_k)
        return self.n_experts * (f * probs.mean(dim=0)).sum()

    def forward(self, x: torch.Tensor):
        """x ``[T, H]`` → ``(weights [T,k], idx [T,k], keep [T,k] bool, capacity, aux_loss | None)``."""
        logits = self._compute_gate_scores(x)
        weights, idx = self._select_experts(logits)
        T = x.shape[0]
        cap = self._compute_expert_capacity(T)
        # position of every (token, choice) inside its expert's queue; earlier choices (k=0) take priority
        onehot = F.one_hot(idx.t().reshape(-1), self.n_experts)            # [k*T, E], k-major order
        pos = (torch.cumsum(onehot, dim=0) - 1) * onehot
        pos = pos.sum(-1).view(self.top_k, T).t()                           # [T, k]
        keep = pos < cap
        loss = None
        if self.training:
            if self.config.use_aux_loss:
                loss = self.config.aux_loss_weight * self.compute_aux_loss(F.softmax(logits, dim=-1), idx)
            if self.config.use_router_z_loss:
                z = self.config.router_z_loss_weight * self.compute_router_z_loss(logits)
                loss = z if loss is None else loss + z
        return weights, idx, keep, cap, loss


class MOELayer(nn.Module):
    """Capacity-based MoE feed-forward: kept (token, expert) pairs are sorted by expert, run through the grouped expert MLP and scattered back with their gate weights; dropped pairs contribute zero."""
    def __init__(self, config: GPTConfig) -> None:
        super().__init__()
        self.router, self.experts = Router(config), MLPExperts(config)

    def forward(self, x: torch.Tensor):
        B, S, H = x.shape
        xt = x.reshape(-1, H)
        weights, idx, keep, _cap, aux = self.router(xt)
        k = idx.shape[1]
        flat_e = idx.reshape(-1)
        flat_keep = keep.reshape(-1)
        # dropped pairs are sent to a virtual expert E so that sorting puts them last
        E = self.router.n_experts
        key = torch.where(flat_keep, flat_e, torch.full_like(flat_e, E))
        order = torch.argsort(key, stable=True)
        token_of = torch.div(order, k, rounding_mode="floor")
        counts = torch.bincount(key, minlength=E + 1)[:E]
        xs = xt.index_select(0, token_of)
        ys = self.experts(xs, counts)           # rows beyond sum(counts) (dropped) come back as zeros
        w = (weights.reshape(-1) * flat_keep).index_select(0, order).to(ys.dtype)
        out = torch.zeros_like(xt).index_add(0, token_of, ys * w.unsqueeze(-1))
        return out.view(B, S, H), aux


class MoEBlock(nn.Module):
    """Pre-norm transformer block whose feed-forward is a ``MOELayer``; returns ``(x, aux_loss)``."""
    def __init__(self, config: GPTConfig) -> None:
        super().__init__()
        self.ln_1, self.ln_2 = LayerNorm(config.n_embd, config.bias), LayerNorm(config.n_embd, config.bias)
        self.attn, self.moe = CausalSelfAttention(config), MOELayer(config)

    def forward(self, x: torch.Tensor):
        x = x + self.attn(self.ln_1(x))
        y, aux = self.moe(self.ln_2(x))
        return x + y, aux


class GPT(nn.Module):
    """The language model: token + position embeddings, dense and MoE blocks as configured, tied LM head; ``forward`` adds the routers' losses to the cross-entropy; ``generate``, ``crop_block_size``, ``estimate_mfu``."""
    def __init__(self, config: GPTConfig) -> None:
        super().__init__()
        self.config = config
        self.transformer = nn.ModuleDict(dict(
            wte=nn.Embedding(config.vocab_size, config.n_embd),
            wpe=nn.Embedding(config.block_size, config.n_embd),
            drop=nn.Dropout(config.dropout),
            blocks=nn.ModuleList([MoEBlock(config) if config.is_moe_layer(i) else Block(config)
                                  for i in range(config.n_layer)]),
            ln_f=LayerNorm(config.n_embd, config.bias)))
        self.lm_head = nn.Linear(config.n_embd, config.vocab_size, bias=False)
        self.transformer.wte.weight = self.lm_head.weight  # weight tying
        self.apply(self._init_weights)
        for n, p in self.named_parameters():
            if n.endswith("c_proj.weight"):
                nn.init.normal_(p, mean=0.0, std=0.02 / math.sqrt(2 * config.n_layer))

    def get_num_params(self, non_embedding: bool = True) -> int:
        n = sum(p.numel() for p in self.parameters())
        return n - self.transformer.wpe.weight.numel() if non_embedding else n

    def _init_weights(self, module: nn.Module) -> None:
        if isinstance(module, nn.Linear):
            nn.init.normal_(module.weight, mean=0.0, std=0.02)
            if module.bias is not None:
                nn.init.zeros_(module.bias)
        elif isinstance(module, nn.Embedding):
            nn.init.normal_(module.weight, mean=0.0, std=0.02)

    def forward(self, idx: torch.Tensor, targets: torch.Tensor | None = None):
        _b, t = idx.shape
        if t > self.config.block_size:
            raise ValueError(f"Sequence length {t} exceeds block size {self.config.block_size}")
        pos = torch.arange(t, dtype=torch.long, device=idx.device)
        x = self.transformer.drop(self.transformer.wte(idx) + self.transformer.wpe(pos))
        total_aux, n_aux = None, 0
        for block in self.transformer.blocks:
            if isinstance(block, MoEBlock):
                x, aux = block(x)
                if aux is not None:
                    total_aux = aux if total_aux is None else total_aux + aux
                    n_aux += 1
            else:
                x = block(x)
        x = self.transformer.ln_f(x)
        if targets is None:
            return self.lm_head(x[:, [-1], :]), None   # inference: last position only
        logits = self.lm_head(x)
        loss = F.cross_entropy(logits.view(-1, logits.size(-1)), targets.view(-1), ignore_index=-1)
        if n_aux > 0:
            loss = loss + total_aux / n_aux
        return logits, loss

    def crop_block_size(self, block_size: int) -> None:
        if block_size > self.config.block_size:
            raise ValueError("cannot grow the block size")
        self.config.block_size = block_size
        self.transformer.wpe.weight = nn.Parameter(self.transformer.wpe.weight[:block_size])

    @torch.no_grad()
    def generate(self, idx: torch.Tensor, max_new_tokens: int, temperature: float = 1.0, top_k: int | None = None):
        for _ in range(max_new_tokens):
            ctx = idx if idx.size(1) <= self.config.block_size else idx[:, -self.config.block_size:]
            logits, _ = self(ctx)
            logits = logits[:, -1, :] / max(temperature, 1e-8)
            if top_k is not None:
                v, _ = torch.topk(logits, min(top_k, logits.size(-1)))
                logits[logits < v[:, [-1]]] = float("-inf")
            idx = torch.cat((idx, torch.multinomial(F.softmax(logits, dim=-1), num_samples=1)), dim=1)
        return idx

    def estimate_mfu(self, fwdbwd_per_iter: int, dt: float, peak_flops: float | None = None) -> float:
        """PaLM-appendix-B estimate ``6N + 12·L·H·Q·T`` per token against the device peak."""
        from ..utils.device import get_theoretical_flops

        cfg = self.config
        N = self.get_num_params()
        L, H, Q, T = cfg.n_layer, cfg.n_head, cfg.n_embd // cfg.n_head, cfg.block_size
        flops_per_iter = (6 * N + 12 * L * H * Q * T) * T * fwdbwd_per_iter
        return (flops_per_iter / dt) / (peak_flops or get_theoretical_flops())


def analyze_moe_usage(model: GPT) -> dict[str, Any]:
    """Which blocks of ``model`` are MoE blocks, how many, and the parameter count."""
    moe_layers = [i for i, b in enumerate(model.transformer.blocks) if isinstance(b, MoEBlock)]
    return {"total_layers": len(model.transformer.blocks), "moe_layers": moe_layers, "moe_layer_count": len(moe_layers),
            "total_parameters": model.get_num_params(), "uses_moe": len(moe_layers) > 0}


def get_moe_layer_info(model: GPT, layer_idx: int) -> dict[str, Any] | None:
    """Expert / router configuration and parameter counts of block ``layer_idx``; ``None`` if it is dense or out of range."""
    if layer_idx >= len(model.transformer.blocks):
        return None
    block = model.transformer.blocks[layer_idx]
    if not isinstance(block, MoEBlock):
        return None
    moe = block.moe
    return {"layer_idx": layer_idx, "n_experts": moe.router.n_experts, "top_k": moe.router.top_k,
            "capacity_factor": moe.router.config.capacity_factor,
            "expert_params": sum(p.numel() for p in moe.experts.parameters()),
            "router_params": sum(p.numel() for p in moe.router.parameters())}
