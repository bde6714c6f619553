"""Miscellaneous helpers (reference: ``scaletorch/utils/misc.py:51-249``)."""
from __future__ import annotations

import fcntl
import os
import random
from typing import Any

import numpy as np
import torch
import torch.distributed as dist

from ..env import ENV_PRINT_LOCK
from .device import get_theoretical_flops, manual_seed_all


def rank_print(*args: Any, is_print_rank: bool = True, **kwargs: Any) -> None:
    """``print`` serialised across processes of a node with an ``flock`` (reference ``misc.py:51-85``)."""
    if not is_print_rank:
        return
    lock_path = os.environ.get(ENV_PRINT_LOCK, "/tmp/scaletorch_b200_print.lock")
    try:
        with open(lock_path, "a") as fh:
            fcntl.flock(fh, fcntl.LOCK_EX)
            try:
                print(*args, **kwargs, flush=True)
            finally:
                fcntl.flock(fh, fcntl.LOCK_UN)
    except OSError:
        print(*args, **kwargs, flush=True)


def set_all_seed(seed: int, deterministic: bool = False, rank_offset: int = 0) -> None:
    """Seeds python / numpy / torch.  ``rank_offset`` lets model-parallel shards initialise
    differently (the reference seeds every rank identically)."""
    if isinstance(seed, bool) or not isinstance(seed, (int, np.integer)):
        raise TypeError(f"seed must be an integer, got {type(seed).__name__}")
    if seed < 0:
        raise ValueError(f"seed must be non-negative, got {seed}")
    seed = int(seed) + int(rank_offset)
    random.seed(seed)
    np.random.seed(seed % (2**32))
    manual_seed_all(seed)
    if deterministic:
        torch.backends.cudnn.deterministic = True
        torch.backends.cudnn.benchmark = False


def to_readable_format(num: int | float, precision: int = 2) -> str:
    """``1234567 -> '1.23M'`` (K / M / B / T; the sign is kept)."""
    if isinstance(num, bool) or not isinstance(num, (int, float, np.integer, np.floating)):
        raise TypeError(f"num must be a number, got {type(num).__name__}")
    if not isinstance(precision, int) or precision < 0:
        raise ValueError(f"precision must be a non-negative integer, got {precision!r}")
    n = float(num)
    for unit, div in (("T", 1e12), ("B", 1e9), ("M", 1e6), ("K", 1e3)):
        if abs(n) >= div:
            return f"{n / div:.{precision}f}{unit}"
    return f"{n:.{precision}f}"


def flops_per_token(num_params: float, num_layers: int, num_heads: int, head_dim: int, seq_len: int) -> float:
    """Training FLOPs per token ``6N + 12·L·heads·head_dim·S`` (reference ``misc.py:171``)."""
    return 6.0 * num_params + 12.0 * num_layers * num_heads * head_dim * seq_len


def get_mfu(tokens_per_second: float, num_params: float, model_config: Any, theoretical_flops: float | None = None,
            seq_len: int | None = None) -> float:
    """Model FLOPs utilisation in percent for *one device's* tokens/s (reference ``misc.py:136-174``)."""
    if tokens_per_second is None or tokens_per_second <= 0 or num_params is None or num_params <= 0:
        return 0.0
    if theoretical_flops is not None and theoretical_flops <= 0:
        return 0.0
    L = model_config.num_hidden_layers
    heads = model_config.num_attention_heads
    hd = getattr(model_config, "head_dim", None) or model_config.hidden_size // heads
    S = seq_len if seq_len is not None else getattr(model_config, "max_position_embeddings", 2048)
    peak = theoretical_flops or get_theoretical_flops()
    return tokens_per_second * flops_per_token(active_params(num_params, model_config), L, heads, hd, S) / peak * 100.0


def active_params(num_params: float, model_config: Any) -> float:
    """Parameters that take part in one token's forward: for MoE models only ``top_k / num_experts`` of the expert
    weights count (the reference's formula uses all of them, which overstates MoE MFU by ~10x)."""
    E = getattr(model_config, "num_experts", 0) or 0
    k = getattr(model_config, "num_experts_per_tok", 0) or 0
    inter = getattr(model_config, "moe_intermediate_size", 0) or 0
    if E <= 0 or k <= 0 or inter <= 0:
        return num_params
    step = getattr(model_config, "decoder_sparse_step", 1) or 1
    mlp_only = getattr(model_config, "mlp_only_layers", None) or []
    n_moe = sum(1 for i in range(model_config.num_hidden_layers) if i not in mlp_only and (i + 1) % step == 0)
    expert_total = n_moe * E * 3 * model_config.hidden_size * inter
    return num_params - expert_total * (1.0 - k / E)


_TP_SHARDED_KEYS = ("q_proj", "k_proj", "v_proj", "qkv_proj", "out_proj", "o_proj", "gate_proj", "up_proj",
                    "gate_up_proj", "down_proj", "embedding", "final_proj")


def is_tp_sharded_name(name: str) -> bool:
    """True for parameters that tensor parallelism shards (by name, like the reference's ``get_num_params``)."""
    return any(k in name for k in _TP_SHARDED_KEYS) and "norm" not in name and "router" not in name and ".experts." not in name


def get_num_params(model: torch.nn.Module) -> int:
    """Global parameter count: TP-sharded tensors × tp, expert tensors × ep, summed over PP stages
    (reference ``misc.py:177-207``)."""
    from ..parallel.process_group import process_group_manager as pgm

    tp = pgm.tp_world_size if pgm else 1
    ep = pgm.ep_world_size if pgm else 1
    seen: set[int] = set()
    total = 0
    for name, p in model.named_parameters():
        if id(p) in seen:
            continue
        seen.add(id(p))
        n = p.numel()
        if ".experts." in name:
            n *= ep
        elif any(k in name for k in _TP_SHARDED_KEYS) and "norm" not in name and "router" not in name:
            n *= tp
        total += n
    if pgm and pgm.pp_world_size > 1:
        t = torch.tensor([total], dtype=torch.int64, device=next(model.parameters()).device)
        if t.device.type == "meta":
            t = torch.tensor([total], dtype=torch.int64)
        dist.all_reduce(t, group=pgm.pp_group)
        total = int(t.item())
    return total


def assert_no_meta_tensors(model: torch.nn.Module) -> None:
    """Raise if any parameter or buffer of ``model`` is still on the meta device (materialisation missed it)."""
    bad = [n for n, p in model.named_parameters() if p.device.type == "meta"]
    bad += [n for n, b in model.named_buffers() if b.device.type == "meta"]
    if bad:
        raise RuntimeError(f"tensors still on the meta device: {bad[:8]}{'...' if len(bad) > 8 else ''}")


def average_loss_across_dp_cp_ranks(loss: float | torch.Tensor | None, device: torch.device, as_tensor: bool = False):
    """Mean of the step loss over the CP×DP(×EP) group on the last PP stage (reference ``misc.py:229-249``).
    ``as_tensor=True`` returns the device scalar without synchronising (for lagged logging)."""
    from ..parallel.process_group import process_group_manager as pgm

    if loss is not None and not isinstance(loss, (int, float, torch.Tensor, np.integer, np.floating)):
        raise TypeError(f"loss must be a number, a tensor or None, got {type(loss).__name__}")
    if isinstance(loss, torch.Tensor):
        t = loss.detach().to(device=device, dtype=torch.float32).reshape(1).clone()
    else:
        t = torch.tensor([0.0 if loss is None else float(loss)], dtype=torch.float32, device=device)
    if pgm and pgm.pp_is_last_stage and pgm.grad_sync_world_size > 1:
        dist.all_reduce(t, group=pgm.grad_sync_group)
        t /= pgm.grad_sync_world_size
    return t if as_tensor else float(t.item())
