"""Slice an unsharded state dict (tp = ep = pp = 1 layout of this framework) onto a parallel model.

Used by the equality tests (parallel run == single-process run) and for re-sharding checkpoints across
layouts, which the reference cannot do (``scaletorch/utils/checkpoint.py:467-560`` requires the same TP/PP
layout on load).
"""
from __future__ import annotations

import re

import torch


def shard_tensor(name: str, full: torch.Tensor, cfg, tp: int, tp_rank: int, ep: int, ep_rank: int) -> torch.Tensor:
    """The slice of the full (local-layout) parameter ``name`` that rank ``(tp_rank, ep_rank)`` owns: head-aligned rows of the fused q|k|v, both halves of gate|up, columns of the row-parallel weights, vocabulary rows, a contiguous block of experts."""
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


# ---------------------------------------------------------------------------------------------- the inverse: merge
def merge_tensor(name: str, pieces: list[torch.Tensor], cfg, tp: int, ep: int) -> torch.Tensor:
    """Inverse of :func:`shard_tensor` for one parameter.  ``pieces`` are the local tensors in rank order: the ``tp``
    tensor-parallel shards for dense parameters, the ``ep`` expert-parallel shards for expert parameters."""
    hd = getattr(cfg, "head_dim", None) or cfg.hidden_size // cfg.num_attention_heads
    nq, nkv = cfg.num_attention_heads, cfg.num_key_value_heads
    if ".experts." in name:
        if len(pieces) != ep:
            raise ValueError(f"{name}: expected {ep} expert-parallel shards, got {len(pieces)}")
        return torch.cat(pieces, dim=0)
    if len(pieces) != tp:
        raise ValueError(f"{name}: expected {tp} tensor-parallel shards, got {len(pieces)}")
    if tp == 1:
        return pieces[0]
    if name.endswith(("qkv_proj.weight", "qkv_proj.bias")):
        parts = [torch.split(t, [nq * hd // tp, nkv * hd // tp, nkv * hd // tp], dim=0) for t in pieces]
        return torch.cat([torch.cat([p[i] for p in parts], dim=0) for i in range(3)], dim=0)      # [q; k; v]
    if name.endswith(("mlp.gate_up_proj.weight", "mlp.gate_up_proj.bias")):
        parts = [torch.chunk(t, 2, dim=0) for t in pieces]
        return torch.cat([torch.cat([p[i] for p in parts], dim=0) for i in range(2)], dim=0)      # [gate; up]
    if name.endswith(("out_proj.weight", "mlp.down_proj.weight")):
        return torch.cat(pieces, dim=1)
    if name in ("embedding.weight", "final_proj.weight"):
        return torch.cat(pieces, dim=0)
    for t in pieces[1:]:                                                   # replicated (norms, router): must agree
        if not torch.equal(t, pieces[0]):
            raise ValueError(f"{name} is replicated over tensor parallel ranks but its copies differ")
    return pieces[0]


def merge_shards(states: dict[tuple[int, int, int], dict[str, torch.Tensor]], cfg, tp: int, pp: int, ep: int,
                 layer_ids_of_stage=None) -> dict[str, torch.Tensor]:
    """Per-coordinate model state dicts ``{(tp_rank, pp_rank, ep_rank): state}`` -> the unsharded (tp = pp = ep = 1)
    state dict of this framework.  Pipeline stages hold locally numbered ``decoder_layers.<i>``; ``layer_ids_of_stage(pp_rank)``
    gives their global indices (default: the contiguous split of ``pipeline_parallel.distribute_layers``).  Dense
    parameters are read from expert-parallel rank 0 (they are replicated over EP), expert parameters from tensor-parallel
    rank 0 (experts are not TP-sharded)."""
    if layer_ids_of_stage is None:
        from ..parallel.pipeline_parallel.pipeline_parallel import distribute_layers

        def layer_ids_of_stage(r):
            return distribute_layers(cfg.num_hidden_layers, pp, r)

    want = {(t, p_, e) for t in range(tp) for p_ in range(pp) for e in range(ep)}
    missing = sorted(want - set(states))
    if missing:
        raise KeyError(f"missing shards for (tp, pp, ep) coordinates {missing[:6]}{'...' if len(missing) > 6 else ''}")
    full: dict[str, torch.Tensor] = {}
    for stage in range(pp):
        ids = layer_ids_of_stage(stage)
        for name in states[(0, stage, 0)]:
            m = re.match(r"decoder_layers\.(\d+)\.(.*)", name)
            gname = f"decoder_layers.{ids[int(m.group(1))]}.{m.group(2)}" if m else name
            if ".experts." in name:
                pieces = [states[(0, stage, e)][name] for e in range(ep)]
            else:
                pieces = [states[(t, stage, 0)][name] for t in range(tp)]
            merged = merge_tensor(gname, pieces, cfg, tp, ep)
            if gname in full:                      # a tied embedding lives on the first AND the last stage
                if not torch.equal(full[gname], merged):
                    raise ValueError(f"{gname} appears on several pipeline stages with different values")
                continue
            full[gname] = merged
    if getattr(cfg, "tie_word_embeddings", False) and "embedding.weight" in full:
        full.setdefault("final_proj.weight", full["embedding.weight"])
    return full
