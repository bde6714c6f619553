"""Zig-zag (load-balanced) sequence partitioning for causal context parallelism.

With the contiguous split of the reference (``data/dataloader.py:222-227``, ``context_parallel.py:162-166``) rank ``r``
attends to ``r + 1`` blocks under a causal mask: the last rank does ``cp`` times the attention work of the first and
everybody waits for it (measured: Qwen3-1.7B cp2 S8192 runs at 115.8 ms/step where a balanced split would need ≈ 85).
Here the sequence is cut into ``2·cp`` chunks and rank ``r`` owns chunks ``r`` and ``2·cp − 1 − r``.  For a peer block
``j``:

* ``j < r``: both local chunks lie after the peer's first chunk and before its second → all local queries attend to the
  *first half* of the peer's K/V, unmasked;
* ``j > r``: only the local second chunk lies after the peer's chunks → the *second half* of the local queries attends
  to all of the peer's K/V, unmasked;
* ``j = r``: ordinary causal attention in local order (local order is position order).

Every rank therefore does one causal block plus ``cp − 1`` half blocks.  Enabled with ``SCALETORCH_B200_CP_ZIGZAG=1``
(the trainer's ``--cp_zigzag``, default on there); the loader, the RoPE positions and both ring implementations read the
same switch.
"""
from __future__ import annotations

import os

import torch

ENV = "SCALETORCH_B200_CP_ZIGZAG"


def enabled() -> bool:
    """Whether the zig-zag split is active (``SCALETORCH_B200_CP_ZIGZAG=1``)."""
    return os.environ.get(ENV, "0") == "1"


def set_enabled(on: bool) -> None:
    """Switch the zig-zag split on or off for the loader, the RoPE positions and both ring implementations."""
    os.environ[ENV] = "1" if on else "0"


def chunk_ranges(seq_len: int, cp: int, rank: int) -> tuple[tuple[int, int], tuple[int, int]]:
    """Global ``[lo, hi)`` ranges of the two chunks owned by ``rank``."""
    if seq_len % (2 * cp) != 0:
        raise ValueError(f"zig-zag context parallelism needs sequence_length ({seq_len}) divisible by 2·cp ({2 * cp})")
    c = seq_len // (2 * cp)
    return (rank * c, (rank + 1) * c), ((2 * cp - 1 - rank) * c, (2 * cp - rank) * c)


def take(x: torch.Tensor, seq_len: int, cp: int, rank: int, dim: int = 1, shift: int = 0) -> torch.Tensor:
    """This rank's two chunks of ``x`` along ``dim`` (``shift = 1`` gives the next-token targets)."""
    (a0, a1), (b0, b1) = chunk_ranges(seq_len, cp, rank)
    return torch.cat([x.narrow(dim, a0 + shift, a1 - a0), x.narrow(dim, b0 + shift, b1 - b0)], dim=dim)


def local_positions(s_local: int, cp: int, rank: int, device=None) -> torch.Tensor:
    """Global token positions of the ``s_local`` local tokens (int64)."""
    c = s_local // 2
    a = torch.arange(rank * c, (rank + 1) * c, device=device)
    b = torch.arange((2 * cp - 1 - rank) * c, (2 * cp - rank) * c, device=device)
    return torch.cat([a, b])


def block_plan(step: int, rank: int, cp: int, s_local: int, causal: bool, zigzag: bool):
    """What rank ``rank`` computes against the K/V block of rank ``(rank − step) mod cp``:
    ``None`` (nothing) or ``(q_lo, k_hi, causal_flag)`` = queries ``[q_lo:]`` × keys ``[:k_hi]``."""
    if step == 0:
        return 0, s_local, causal
    if not causal:
        return 0, s_local, False
    src = (rank - step) % cp
    if not zigzag:
        return (0, s_local, False) if src < rank else None
    c = s_local // 2
    return (0, c, False) if src < rank else (c, s_local, False)
