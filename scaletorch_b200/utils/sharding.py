"""Slice an unsharded state dict (tp = ep = pp = 1 layout of this framework) onto a parallel model.

Used by the equality tests (parallel run == single-process run) and for re-sharding checkpoints across
layouts, which the reference cannot do (``scaletorch/utils/checkpoint.py:467-560`` requires the same TP/PP
layout on load).
"""
from __future__ import annotations

import re

import torch


def shard_tensor(name: str, full: torch.Tensor, cfg, tp: int, tp_rank: int, ep: int, ep_rank: int) -> torch.Tensor:
    hd = getattr(cfg, "head_dim", None) or cfg.hidden_size // cfg.num_attention_heads
    nq, nkv = cfg.num_attention_heads, cfg.num_key_value_heads

    def rows(t, n_parts=tp, r=tp_rank):
        n = t.shape[0] // n_parts
        return t[r * n:(r + 1) * n]

    if name.endswith(("qkv_proj.weight", "qkv_proj.bias")):
        q, k, v = torch.split(full, [nq * hd, nkv * hd, nkv * hd], dim=0)
        return torch.cat([rows(q), rows(k), rows(v)], dim=0)
    if name.endswith(("mlp.gate_up_proj.weight", "mlp.gate_up_proj.bias")):
        g, u = torch.chunk(full, 2, dim=0)
        return torch.cat([rows(g), rows(u)], dim=0)
    if name.endswith(("out_proj.weight", "mlp.down_proj.weight")):
        n = full.shape[1] // tp
        return full[:, tp_rank * n:(tp_rank + 1) * n]
    if name in ("embedding.weight", "final_proj.weight"):
        return rows(full)
    if ".experts." in name:
        return rows(full, ep, ep_rank)
    return full


def load_from_unsharded(par_model: torch.nn.Module, full_state: dict[str, torch.Tensor], cfg, tp: int = 1,
                        tp_rank: int = 0, ep: int = 1, ep_rank: int = 0, layer_ids: list[int] | None = None) -> None:
    """Copies the right slice of every tensor of ``full_state`` into ``par_model`` (in place)."""
    with torch.no_grad():
        for name, p in par_model.named_parameters():
            src_name = name
            if layer_ids is not None:
                m = re.match(r"decoder_layers\.(\d+)\.(.*)", name)
                if m:
                    src_name = f"decoder_layers.{layer_ids[int(m.group(1))]}.{m.group(2)}"
            if src_name not in full_state:
                if src_name == "final_proj.weight" and "embedding.weight" in full_state:
                    src_name = "embedding.weight"      # tied embeddings
                else:
                    raise KeyError(f"{src_name} missing from the unsharded state dict")
            piece = shard_tensor(src_name, full_state[src_name], cfg, tp, tp_rank, ep, ep_rank)
            if tuple(piece.shape) != tuple(p.shape):
                raise ValueError(f"{name}: shard shape {tuple(piece.shape)} != parameter shape {tuple(p.shape)}")
            p.copy_(piece.to(p.dtype))
