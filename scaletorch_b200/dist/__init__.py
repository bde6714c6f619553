"""Distributed communication primitives over ``torch.distributed`` (NCCL on B200, gloo on CPU).

Same 28 exported names as the reference (``scaletorch/dist/__init__.py:56-85``).  These wrappers are
the plumbing / baseline path; the hot collectives of TP/SP/CP/EP/DP run as fused kernels over
symmetric memory (``scaletorch_b200.runtime.symm`` and ``scaletorch_b200.parallel``).
"""
from ._reduce_op import _get_reduce_op  # noqa: F401
from .collective_ops import (all_gather, all_reduce, all_reduce_dict, all_reduce_params, all_to_all, broadcast,
                             gather, reduce, reduce_scatter, scatter, sync_random_seed)
from .gather_utils import collect_results, collect_results_cpu, collect_results_gpu
from .object_ops import all_gather_object, broadcast_object_list, gather_object
from .p2p_ops import P2POp, batch_isend_irecv
from .utils import (barrier, cleanup_dist, destroy_group, get_rank, get_world_size, init_dist, is_distributed,
                    new_group)

__all__ = [
    "_get_reduce_op", "all_gather", "all_gather_object", "all_reduce", "all_reduce_dict", "all_reduce_params",
    "all_to_all", "barrier", "batch_isend_irecv", "broadcast", "broadcast_object_list", "cleanup_dist",
    "collect_results", "collect_results_cpu", "collect_results_gpu", "destroy_group", "gather", "gather_object",
    "get_rank", "get_world_size", "init_dist", "is_distributed", "new_group", "P2POp", "reduce",
    "reduce_scatter", "scatter", "sync_random_seed",
]
