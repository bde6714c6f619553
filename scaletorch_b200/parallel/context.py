"""``ParallelContext``: the explicit, resolved-once description of how a model is parallelised.

The reference steers models through environment variables read inside module constructors
(``CONTEXT_PARALLEL``, ``SEQUENCE_PARALLEL``, ``FLASH_ATTEN`` -- ``scaletorch/models/model_qwen3.py:182-186``,
``scaletorch/models/llama.py:314-318``), which makes the result depend on call order
(SURVEY §2.7 #6).  Here the env vars are still honoured as *inputs*, but they are folded into this
object exactly once and the object is handed to the model.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Any

from ..env import ENV_FLASH_ATTENTION, ENV_FUSED_COMM, ENV_SEQUENCE_PARALLEL
from .process_group import process_group_manager as pgm


@dataclass
class ParallelContext:
    """Sizes, ranks and groups of the TP / CP / EP / PP / DP axes plus the resolved switches (sequence parallel, flash attention, fused collectives) that a model is built with."""
    tp_size: int = 1
    tp_rank: int = 0
    tp_group: Any = None
    cp_size: int = 1
    cp_rank: int = 0
    cp_group: Any = None
    ep_size: int = 1
    ep_rank: int = 0
    ep_group: Any = None
    pp_size: int = 1
    pp_rank: int = 0
    dp_size: int = 1
    dp_rank: int = 0
    sequence_parallel: bool = False   # Megatron-SP around the norm regions (needs tp_size > 1)
    use_flash_attention: bool = True  # False -> eager oracle attention (debug)
    fused_comm: bool = True           # symmetric-memory fused collectives when available

    @property
    def sp(self) -> bool:
        return self.sequence_parallel and self.tp_size > 1

    @classmethod
    def from_pgm(cls, sequence_parallel: bool | None = None, use_flash_attention: bool | None = None,
                 fused_comm: bool | None = None) -> "ParallelContext":
        if sequence_parallel is None:
            sequence_parallel = os.environ.get(ENV_SEQUENCE_PARALLEL, "0") == "1"
        if use_flash_attention is None:
            use_flash_attention = os.environ.get(ENV_FLASH_ATTENTION, "1") == "1"
        if fused_comm is None:
            fused_comm = os.environ.get(ENV_FUSED_COMM, "1") == "1"
        if not pgm:
            return cls(sequence_parallel=False, use_flash_attention=use_flash_attention, fused_comm=fused_comm)
        return cls(
            tp_size=pgm.tp_world_size, tp_rank=pgm.tp_rank, tp_group=pgm.tp_group,
            cp_size=pgm.cp_world_size, cp_rank=pgm.cp_rank, cp_group=pgm.cp_group,
            ep_size=pgm.ep_world_size, ep_rank=pgm.ep_rank, ep_group=pgm.ep_group,
            pp_size=pgm.pp_world_size, pp_rank=pgm.pp_rank,
            dp_size=pgm.dp_world_size, dp_rank=pgm.dp_rank,
            sequence_parallel=bool(sequence_parallel) and pgm.tp_world_size > 1,
            use_flash_attention=use_flash_attention, fused_comm=fused_comm,
        )


def single_device_context(**kw) -> ParallelContext:
    """A context with every axis of size 1 (keyword arguments override fields)."""
    return ParallelContext(**kw)
