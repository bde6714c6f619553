"""Ring communicator for context parallelism (reference:
``scaletorch/parallel/context_parallel/cp_comms.py:38-232``).

State machine ``send_recv* → commit → wait`` with the same misuse assertions as the reference
(commit twice / wait before commit).  This is the NCCL / gloo baseline transport; the B200 ring
attention kernel reads the neighbour's K/V straight from peer HBM instead (``parallel/fused.py``).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from ...env import ENV_VERBOSE
from ..process_group import process_group_manager as pgm

VERBOSE = os.environ.get(ENV_VERBOSE, "0") == "1"


class ContextCommunicator:
    """Send-to-next / receive-from-previous on the CP ring as a small state machine: ``send_recv`` queues pairs, ``commit`` launches them as one batch, ``wait`` completes them; misuse (commit twice, wait before commit) raises."""
    def __init__(self, msg: str = "", group=None, send_rank: int | None = None, recv_rank: int | None = None):
        self.msg = msg
        self._ops: list = []
        self._reqs: list | None = None
        self._pending: list = []
        if group is None and pgm:
            group, send_rank, recv_rank = pgm.cp_group, pgm.cp_send_rank, pgm.cp_recv_rank
        self.group = group
        self.send_rank, self.recv_rank = send_rank, recv_rank
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.world_size = dist.get_world_size(group) if group is not None else 1

    def send_recv(self, tensor_to_send: torch.Tensor, recv_tensor: torch.Tensor | None = None) -> torch.Tensor:
        """Queues ``send(tensor → next rank)`` + ``recv(← previous rank)``; returns the receive buffer."""
        if self._reqs is not None:
            raise RuntimeError("send_recv() after commit(): call wait() first")
        if recv_tensor is None:
            recv_tensor = torch.empty(tensor_to_send.shape, dtype=tensor_to_send.dtype, device=tensor_to_send.device)
        send = tensor_to_send.contiguous()
        if self.world_size > 1:
            self._ops.append(dist.P2POp(dist.isend, send, self.send_rank, group=self.group))
            self._ops.append(dist.P2POp(dist.irecv, recv_tensor, self.recv_rank, group=self.group))
            self._pending.append(send)  # keep the staging tensor alive until wait()
        else:
            recv_tensor.copy_(send)
        return recv_tensor

    def commit(self) -> None:
        if self._reqs is not None:
            raise RuntimeError("commit called twice")
        if VERBOSE:
            print(f"[cp {self.msg}] rank {self.rank}: commit {len(self._ops)} ops → {self.send_rank}, ← {self.recv_rank}",
                  flush=True)
        self._reqs = dist.batch_isend_irecv(self._ops) if self._ops else []

    def wait(self) -> None:
        if self._reqs is None:
            raise RuntimeError("wait called before commit")
        for r in self._reqs:
            r.wait()
        self._reqs = None
        self._ops = []
        self._pending = []
