"""Model and optimiser construction (reference: ``scaletorch/trainer/model_builder.py:33-184``).

Pipeline: resolve config → build on the meta device from a :class:`ParallelContext` (already TP/SP/CP/EP
aware) → pipeline-parallel wrap → materialise (random init or kept HF weights) → cast/move → flat
parameter store → data-parallel wrap.  ``apply_context_parallel`` runs *before* anything depends on the
attention backend (the reference sets the env var after the model is built).
"""
from __future__ import annotations

import time

import torch

from ..models.llama import Llama
from ..models.model_qwen3 import Qwen3
from ..models.model_qwen3_moe import Qwen3MoE
from ..parallel.context import ParallelContext
from ..parallel.data_parallel.data_parallel import DataParallelBucket, get_store
from ..parallel.process_group import process_group_manager as pgm
from ..utils.checkpoint import init_model_with_dematerialized_weights, init_model_with_materialized_weights
from ..utils.logger_utils import get_logger
from ..utils.misc import rank_print, set_all_seed
from .config import resolve_model_config

logger = get_logger(__name__)

MODEL_CLASSES = {"llama": Llama, "qwen3": Qwen3, "qwen3_moe": Qwen3MoE}


def model_class_for(model_config):
    """Model class for an HF config by its ``model_type`` (``llama``, ``qwen3``, ``qwen3_moe``; unknown types fall back to Llama)."""
    return MODEL_CLASSES.get(getattr(model_config, "model_type", "llama"), Llama)


def create_model(config, dtype: torch.dtype, device: torch.device):
    """Returns ``(model, model_config)``; ``model`` is wrapped for PP / DP as the layout requires."""
    t0 = time.time()
    model_config = resolve_model_config(config.model_name_or_path)
    if getattr(config, "ep_capacity_factor", None) is not None:
        model_config.ep_capacity_factor = float(config.ep_capacity_factor)
    for k in ("num_hidden_layers", "num_attention_heads", "num_key_value_heads"):
        v = getattr(config, k, None)
        if v is not None and v != getattr(model_config, k, None):
            setattr(model_config, k, v)
    ctx = ParallelContext.from_pgm(sequence_parallel=getattr(config, "sequence_parallel", None) or None,
                                   use_flash_attention=getattr(config, "use_flash_attention", True),
                                   fused_comm=getattr(config, "fused_comm", True))
    # model-parallel shards must not start from identical values (reference seeds all ranks alike, §2.7 #14)
    if pgm:
        set_all_seed(config.seed, rank_offset=1000 * pgm.tp_rank + 100_000 * pgm.pp_rank + 10_000_000 * pgm.ep_rank)
    with init_model_with_dematerialized_weights():
        model = model_class_for(model_config)(model_config, ctx)
        if pgm and pgm.pp_world_size > 1:
            from ..parallel.pipeline_parallel.pipeline_parallel import PipelineParallel

            model = PipelineParallel(model, model_config)
    model = init_model_with_materialized_weights(
        model, model_config, save_dir=config.model_name_or_path,
        load_weights=getattr(config, "load_pretrained_weights", False))
    model.to(dtype).to(device)
    if pgm:
        sync_replicated_parameters(model)
    if pgm and pgm.tp_world_size > 1:
        from ..parallel.tensor_parallel.tensor_parallel import mark_tp_partial_params

        mark_tp_partial_params(model, pgm.tp_group, ctx.sp)
    grad_sync = pgm.grad_sync_world_size if pgm else 1
    # flat parameter / main-grad buffers (also on a single GPU); in symmetric memory when gradients are
    # all-reduced over NVLink by our own kernel
    store = get_store(model, pgm.grad_sync_group if (pgm and grad_sync > 1) else None, getattr(config, "fused_comm", True))
    if pgm and pgm.tp_world_size > 1 and store is not None:
        store.tp_group = pgm.tp_group
    if pgm and (grad_sync > 1 or pgm.expert_grad_sync_world_size > 1):
        wire = torch.bfloat16 if getattr(config, "grad_wire_dtype", "float32") == "bfloat16" else torch.float32
        model = DataParallelBucket(model, wire_dtype=wire)
    rank_print(f"init model parallel time: {time.time() - t0:.2f}s",
               is_print_rank=(not pgm) or pgm.global_rank == 0)
    return model, model_config


@torch.no_grad()
def sync_replicated_parameters(model: torch.nn.Module) -> int:
    """Makes every *replicated* parameter start bit-equal on the ranks that hold a copy of it.

    The per-rank seed offset above exists so that *shards* differ; it must not leak into tensors that an axis only
    replicates, because their gradients are summed/averaged over that axis on the assumption that the copies are one
    tensor.  After materialising: tensors TP does not shard (norms, router, stacked experts) are broadcast from TP
    rank 0; everything but the experts is broadcast from EP rank 0; with tied embeddings split over pipeline stages
    the first stage's embedding is sent to the last stage's LM-head copy.  Returns the number of tensors broadcast.
    """
    import torch.distributed as dist

    from ..utils.misc import is_tp_sharded_name

    if not (dist.is_available() and dist.is_initialized()):
        return 0
    n = 0
    named = [(k, p) for k, p in model.named_parameters()]
    if pgm.tp_world_size > 1:
        src = dist.get_global_rank(pgm.tp_group, 0)
        for k, p in named:
            if not is_tp_sharded_name(k):
                dist.broadcast(p.data, src=src, group=pgm.tp_group)
                n += 1
    if pgm.ep_world_size > 1:
        src = dist.get_global_rank(pgm.ep_group, 0)
        for k, p in named:
            if ".experts." not in k:
                dist.broadcast(p.data, src=src, group=pgm.ep_group)
                n += 1
    raw = unwrap(model)
    if getattr(raw, "tie_word_embeddings", False) and pgm.pp_world_size > 1:
        from ..parallel.pipeline_parallel.pipeline_parallel import _first_last_group

        group = _first_last_group()          # created by every rank in PipelineParallel.__init__
        if group is not None and getattr(raw, "tied_across_stages", False):
            w = raw.embedding.weight if raw.is_first else raw.final_proj.weight
            dist.broadcast(w.data, src=dist.get_global_rank(group, 0), group=group)
            n += 1
    return n


def unwrap(model):
    """Strip ``.module`` wrappers (data parallel, pipeline stage) down to the bare model."""
    while hasattr(model, "module") and isinstance(model.module, torch.nn.Module):
        model = model.module
    return model


def create_optimizer(model: torch.nn.Module, config, device: torch.device):
    """adamw (fused flat kernel) | adam | sgd (momentum .9) | lamb → adamw (reference ``model_builder.py:103-162``)."""
    opt = getattr(config, "optimizer_type", "adamw")
    wd = getattr(config, "weight_decay", 0.0)
    betas = tuple(getattr(config, "betas", (0.9, 0.999)))
    lr = getattr(config, "learning_rate", 1e-3)
    raw = unwrap(model)
    if opt == "lamb":
        logger.warning("LAMB is not available; falling back to AdamW")
        opt = "adamw"
    if opt == "adamw" and getattr(config, "use_fused_adam", True):
        from .optimizer import FusedAdamW

        store = get_store(raw)
        norm_group = pgm.mp_group if (pgm and len(pgm.mp_group_ids) > 1) else None
        kw = dict(lr=lr, betas=betas, weight_decay=wd, max_grad_norm=getattr(config, "max_grad_norm", None),
                  norm_group=norm_group)
        bm = getattr(model, "bucket_manager", None)
        if (getattr(store, "param_window", None) is not None and bm is not None and pgm and pgm.pp_world_size == 1
                and pgm.grad_sync_world_size > 1 and getattr(config, "zero_stage", 0) != -1):
            # gradient buffer and parameters are in symmetric memory: ZeRO-1 on the peer-memory kernels (reduce-scatter
            # during backward, sharded AdamW that writes the new parameters into every rank's copy)
            from .zero import PeerShardedAdamW

            return PeerShardedAdamW(store, bm, pgm.grad_sync_group, **kw)
        if getattr(config, "zero_stage", 0) >= 1 and pgm and pgm.grad_sync_world_size > 1:
            from .zero import ZeroFusedAdamW

            return ZeroFusedAdamW(store, dense_group=pgm.grad_sync_group,
                                  expert_group=pgm.expert_grad_sync_group if pgm.expert_grad_sync_world_size > 1 else None,
                                  **kw)
        return FusedAdamW(store, **kw)
    params = [p for p in raw.parameters() if p.requires_grad]
    if opt == "adamw":
        return torch.optim.AdamW(params, lr=lr, betas=betas, weight_decay=wd)
    if opt == "adam":
        return torch.optim.Adam(params, lr=lr, betas=betas, weight_decay=wd)
    if opt == "sgd":
        return torch.optim.SGD(params, lr=lr, momentum=0.9, weight_decay=wd)
    raise ValueError(f"unknown optimizer_type {opt}")


def get_tensor_shapes(config, model_config) -> tuple[int, int, int]:
    """Shape of the activation exchanged between pipeline stages: ``(mbs, S/cp, hidden)``."""
    s_local = config.sequence_length // config.context_parallel_size
    if getattr(config, "sequence_parallel", False) and config.tensor_parallel_size > 1:
        s_local //= config.tensor_parallel_size   # the boundary tensor is the TP rank's token shard
    return (config.micro_batch_size, s_local, model_config.hidden_size)
