"""String → ``ReduceOp`` mapping (reference: ``scaletorch/dist/_reduce_op.py:7-25``).  ``'mean'`` maps to
SUM; callers divide by the group size afterwards (same convention as the reference)."""
import torch.distributed as dist

_OPS = {
    "sum": dist.ReduceOp.SUM,
    "mean": dist.ReduceOp.SUM,
    "product": dist.ReduceOp.PRODUCT,
    "min": dist.ReduceOp.MIN,
    "max": dist.ReduceOp.MAX,
    "band": dist.ReduceOp.BAND,
    "bor": dist.ReduceOp.BOR,
    "bxor": dist.ReduceOp.BXOR,
}


def _get_reduce_op(name: str):
    key = name.lower()
    if key not in _OPS:
        raise ValueError(f"reduce op should be one of {sorted(_OPS)}, but got {name}")
    return _OPS[key]
