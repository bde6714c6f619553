"""Sequence-parallel region transitions (reference: ``scaletorch/parallel/sequence_parallel/sp_comms.py``).

The two autograd functions of the reference, exposed under their original names.  Activations are
token-major ``[T, H]`` here, so the sequence dimension is dim 0 and the gather needs no ``torch.cat``.
"""
from ..tensor_parallel.tp_comms import (_AllGatherDim0 as AllGatherFromSequenceParallelRegion,
                                        _ReduceScatterDim0 as ReduceScatterToSequenceParallelRegion,
                                        gather_from_sequence_parallel_region,
                                        reduce_scatter_to_sequence_parallel_region, split_along_dim0)

__all__ = ["AllGatherFromSequenceParallelRegion", "ReduceScatterToSequenceParallelRegion",
           "gather_from_sequence_parallel_region", "reduce_scatter_to_sequence_parallel_region",
           "split_along_dim0"]
