"""Point-to-point helpers (reference: ``scaletorch/dist/p2p_ops.py:12-42``)."""
from __future__ import annotations

import torch.distributed as torch_dist
from torch import Tensor


def isend(tensor: Tensor, dst: int, group=None, tag: int = 0):
    """Non-blocking send of ``tensor`` to GLOBAL rank ``dst``; returns the work handle."""
    return torch_dist.isend(tensor, dst, group=group, tag=tag)


def irecv(tensor: Tensor, src: int, group=None, tag: int = 0):
    """Non-blocking receive into ``tensor`` from GLOBAL rank ``src``; returns the work handle."""
    return torch_dist.irecv(tensor, src, group=group, tag=tag)


def P2POp(op, tensor: Tensor, peer: int, group=None, tag: int = 0):
    """``op`` is ``torch.distributed.isend`` / ``irecv`` (or this module's wrappers)."""
    if op is isend:
        op = torch_dist.isend
    elif op is irecv:
        op = torch_dist.irecv
    return torch_dist.P2POp(op, tensor, peer, group, tag)


def batch_isend_irecv(p2p_ops: list) -> list:
    """Launch a list of ``P2POp`` as one batch (one NCCL group call) and return their work handles."""
    if not p2p_ops:
        return []
    return torch_dist.batch_isend_irecv(p2p_ops)
