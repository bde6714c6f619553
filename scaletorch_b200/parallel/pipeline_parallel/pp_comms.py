"""Pipeline-parallel point-to-point communication (reference:
``scaletorch/parallel/pipeline_parallel/pp_comms.py:17-285``).

Same operation vocabulary (``recv_forward / send_forward / recv_backward / send_backward`` and the fused
``send_fwd_recv_bwd / send_bwd_recv_fwd``).  Stays on NCCL p2p by design (BASELINE north star), but:
receives are *pre-posted* with ``batch_isend_irecv`` and only waited on when the tensor is consumed, so a
stage's compute overlaps the transfer instead of blocking on ``dist.recv`` on the compute stream
(reference ``pp_comms.py:182-186,264-266``).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from ...env import ENV_VERBOSE
from ..process_group import process_group_manager as pgm

STEP = 0
VERBOSE = os.environ.get(ENV_VERBOSE, "0") == "1"

_OPS = {"recv_forward", "send_forward", "recv_backward", "send_backward"}


class PendingRecv:
    """A posted receive; ``wait()`` returns the tensor."""

    def __init__(self, tensor: torch.Tensor, reqs: list):
        self.tensor, self.reqs = tensor, reqs

    def wait(self) -> torch.Tensor:
        for r in self.reqs:
            r.wait()
        self.reqs = []
        return self.tensor


def _log(op: str, peer: int | None) -> None:
    global STEP
    if VERBOSE:
        print(f"[pp step {STEP}] rank {pgm.global_rank} (stage {pgm.pp_rank}): {op} peer={peer}", flush=True)
    STEP += 1


def post_recv(operation: str, shapes, device, dtype) -> PendingRecv | None:
    """Posts the receive for ``recv_forward`` / ``recv_backward`` without waiting."""
    if operation == "recv_forward":
        if pgm.pp_is_first_stage:
            return None
        src = pgm.pp_prev_rank
    elif operation == "recv_backward":
        if pgm.pp_is_last_stage:
            return None
        src = pgm.pp_next_rank
    else:
        raise ValueError(f"Invalid recv operation: {operation}")
    t = torch.empty(shapes, dtype=dtype, device=device, requires_grad=(operation == "recv_forward"))
    reqs = dist.batch_isend_irecv([dist.P2POp(dist.irecv, t, src)])
    _log(operation, src)
    return PendingRecv(t, reqs)


def pipeline_communicate(operation: str, device=None, dtype=None, tensor: torch.Tensor | None = None, shapes=None):
    """Blocking single-direction transfer (reference API).  Returns the received tensor for recv ops."""
    if operation not in _OPS:
        raise ValueError(f"Invalid operation: {operation}. Must be one of {sorted(_OPS)}")
    if operation.startswith("recv"):
        if shapes is None:
            raise ValueError("shapes must be provided for receive operations")
        pend = post_recv(operation, shapes, device, dtype)
        return None if pend is None else pend.wait()
    if operation == "send_forward":
        if pgm.pp_is_last_stage:
            return None
        dst = pgm.pp_next_rank
    else:
        if pgm.pp_is_first_stage:
            return None
        dst = pgm.pp_prev_rank
    if tensor is None:
        raise ValueError("tensor must be provided for send operations")
    reqs = dist.batch_isend_irecv([dist.P2POp(dist.isend, tensor.contiguous(), dst)])
    _log(operation, dst)
    for r in reqs:
        r.wait()
    return None


def bidirectional_pipeline_communicate(operation: str, send_tensor: torch.Tensor, recv_shapes, device, dtype):
    """Fused send+recv with one neighbour (``send_fwd_recv_bwd`` | ``send_bwd_recv_fwd``)."""
    if operation not in ("send_fwd_recv_bwd", "send_bwd_recv_fwd"):
        raise ValueError(f"Invalid operation: {operation}")
    is_fwd = operation == "send_fwd_recv_bwd"
    if (is_fwd and pgm.pp_is_last_stage) or (not is_fwd and pgm.pp_is_first_stage):
        return None
    peer = pgm.pp_next_rank if is_fwd else pgm.pp_prev_rank
    recv = torch.empty(recv_shapes, dtype=dtype, device=device, requires_grad=not is_fwd)
    reqs = dist.batch_isend_irecv([dist.P2POp(dist.isend, send_tensor.contiguous(), peer),
                                   dist.P2POp(dist.irecv, recv, peer)])
    _log(operation, peer)
    for r in reqs:
        r.wait()
    return recv
