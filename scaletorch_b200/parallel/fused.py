"""Dispatch layer for the fused compute+collective kernels over symmetric memory.

Each ``*_available`` predicate answers "can the fused sm_100a kernel serve this call?" (CUDA bf16
tensors, NCCL group fully inside one NVSwitch domain, symmetric-memory runtime initialised, shapes
supported).  When it returns False the callers in ``tensor_parallel`` / ``data_parallel`` /
``expert_parallel`` / ``context_parallel`` run the c10d path, which is also the measured baseline.
"""
from __future__ import annotations

import torch

from ..runtime import symm


def _ok(t: torch.Tensor, group) -> bool:
    return t.is_cuda and t.dtype == torch.bfloat16 and symm.group_ready(group)


def ag_gemm_available(x2d: torch.Tensor, group) -> bool:
    return _ok(x2d, group) and symm.has_kernel("stb_ag_gemm_bf16")


def gemm_rs_available(x2d: torch.Tensor, group) -> bool:
    return _ok(x2d, group) and symm.has_kernel("stb_gemm_rs_bf16")


def allreduce_available(t: torch.Tensor, group) -> bool:
    return _ok(t, group) and symm.has_kernel("stb_allreduce_bf16")


def all_gather_gemm(x2d, weight, group, b_mn: bool = False):
    from ..runtime import fused_ops

    return fused_ops.all_gather_gemm(x2d, weight, group, b_mn=b_mn)


def gemm_reduce_scatter(x2d, weight, group, b_mn: bool = False):
    from ..runtime import fused_ops

    return fused_ops.gemm_reduce_scatter(x2d, weight, group, b_mn=b_mn)


def all_reduce_(t, group):
    from ..runtime import fused_ops

    return fused_ops.all_reduce_(t, group)
