"""Dispatch layer for the fused compute+collective kernels over symmetric memory.

Each ``*_available`` predicate answers "can the fused sm_100a kernel serve this call?" (CUDA bf16 tensors, an
NCCL group inside one NVSwitch domain, shapes the kernels support, window large enough).  When it returns
False the callers in ``tensor_parallel`` / ``data_parallel`` run the c10d path, which is also the measured
baseline (``--fused_comm False`` or ``SCALETORCH_B200_FUSED_COMM=0`` forces it).
"""
from __future__ import annotations

import torch

from ..runtime import symm


def _ok(t: torch.Tensor, group) -> bool:
    return t.is_cuda and t.dtype == torch.bfloat16 and t.is_contiguous() and symm.group_ready(group)


def _world(group) -> int:
    import torch.distributed as dist

    return dist.get_world_size(group)


def _policy(op: str, world: int) -> bool:
    """Measured choice between the fused kernel and GEMM + c10d (profiles/fused_collectives_r2.md, B200, NVLink 5).

    The fused kernels pay with SMs (copy CTAs), a staging copy or a device barrier; NCCL pays with a separate launch
    and un-overlapped link time.  On a 2-GPU group the link time is 5-10 % of the GEMM and GEMM + NCCL wins in the model
    (Qwen3-8B TP2+SP S4096: 133.7 ms c10d, 135.5 ms with only GEMM->reduce-scatter fused, 150.1 ms with everything fused);
    from 4 ranks on the transfers are a third of the GEMM and GEMM->reduce-scatter / GEMM->all-reduce win (1.13x / 1.24x at
    8 ranks), all-gather->GEMM ties.  ``SCALETORCH_B200_FUSED_TP=all|none|auto`` overrides."""
    import os

    mode = os.environ.get("SCALETORCH_B200_FUSED_TP", "auto")
    if mode == "all":
        return True
    if mode == "none":
        return False
    return world >= 4


def ag_gemm_available(x_shard: torch.Tensor, group) -> bool:
    """x_shard ``[T/world, K]``: every rank's row block must be a multiple of the 128-row tile."""
    return (_ok(x_shard, group) and symm.has_kernel("stb_ag_gemm_bf16") and x_shard.shape[0] % 128 == 0
            and x_shard.numel() * 2 <= symm.region_bytes(group) and _world(group) <= 8
            and _policy("ag_gemm", _world(group)))


def gemm_rs_available(x_full: torch.Tensor, group) -> bool:
    """x_full ``[T, K]`` with ``T / world`` a multiple of 128."""
    if not _ok(x_full, group):
        return False
    w = _world(group)
    return (symm.has_kernel("stb_gemm_rs_bf16") and x_full.shape[0] % (128 * w) == 0 and w <= 8
            and _policy("gemm_rs", w))


def gemm_ar_available(x: torch.Tensor, n_out: int, group) -> bool:
    """Whether GEMM -> all-reduce through the window applies to ``x @ W`` with ``n_out`` output features (fits the region, 16-byte multiple, policy allows it)."""
    return (_ok(x, group) and symm.has_kernel("stb_allreduce_oneshot_bf16")
            and x.shape[0] * n_out * 2 <= symm.region_bytes(group) and (x.shape[0] * n_out) % 8 == 0
            and _policy("gemm_ar", _world(group)))


def allreduce_available(t: torch.Tensor, group) -> bool:
    """Whether the peer-memory all-reduce applies to ``t`` (bf16 CUDA, multiple of 8 elements, fits the region)."""
    return (_ok(t, group) and symm.has_kernel("stb_allreduce_oneshot_bf16") and t.numel() % 8 == 0
            and t.numel() * 2 <= symm.region_bytes(group))


def all_gather_gemm(x2d, weight, group, b_mn: bool = False):
    """``(x_full @ W.T, x_full)`` from this rank's row shard: the GEMM consumes peer shards as copy CTAs land them (``runtime.fused_ops.all_gather_gemm``)."""
    from ..runtime import fused_ops

    return fused_ops.all_gather_gemm(x2d, weight, group, b_mn=b_mn)


def gemm_reduce_scatter(x2d, weight, group, b_mn: bool = False):
    """This rank's row block of ``sum_over_ranks(x @ W.T)``: the GEMM epilogue reduce-adds each tile into its owner's buffer (``runtime.fused_ops.gemm_reduce_scatter``)."""
    from ..runtime import fused_ops

    return fused_ops.gemm_reduce_scatter(x2d, weight, group, b_mn=b_mn)


def gemm_all_reduce(x2d, weight, group, b_mn: bool = False):
    """``sum_over_ranks(x @ W.T)`` on every rank: GEMM into the window, then a one-shot / NVLS all-reduce kernel (``runtime.fused_ops.gemm_all_reduce``)."""
    from ..runtime import fused_ops

    return fused_ops.gemm_all_reduce(x2d, weight, group, b_mn=b_mn)


def all_reduce_(t, group):
    """Sum a bf16 tensor over the group through the symmetric window; returns a new tensor."""
    from ..runtime import fused_ops

    return fused_ops.all_reduce_(t, group)
