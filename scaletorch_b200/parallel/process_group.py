"""``ProcessGroupManager`` over the 5-D ``[DP, PP, CP, EP, TP]`` rank grid.

Public surface follows the reference (``scaletorch/parallel/process_group.py:22-415``): TP is the
fastest-varying axis and DP the slowest, ``world_size == tp·cp·pp·dp·ep``, seven group families
(TP, CP, PP, EP, DP, CP×DP, PP×DP) and the same attribute names.  Additions for this framework:

* ``grad_sync_group`` / ``expert_grad_sync_group``: when EP ranks consume *distinct* data (the
  default here) dense gradients are reduced over DP×CP×EP and expert gradients over
  DP×CP; with ``ep_replicate_data=True`` (reference semantics: every EP rank sees the same tokens)
  both reduce over DP×CP.
* ``mp_group``: all ranks of one model replica (TP×PP×(EP)) -- the set over which a global gradient
  norm must be reduced (the reference clips with rank-local norms).
* group construction derives every family from the grid tensor with one generic routine instead of
  seven hand-written loops.
"""
from __future__ import annotations

import os
from itertools import product
from typing import Any

import torch

from ..dist import get_rank, get_world_size, is_distributed, new_group

_AXES = ("dp", "pp", "cp", "ep", "tp")


class ProcessGroupManager:
    """The ``[DP, PP, CP, EP, TP]`` rank grid (TP fastest): this rank's coordinates, one process group per axis plus the CP x DP, PP x DP, gradient-sync and model-parallel groups, ring / pipeline neighbours, data rank for the loader."""
    def __init__(self, tp_size: int, cp_size: int, pp_size: int, dp_size: int, ep_size: int = 1,
                 ep_replicate_data: bool = False) -> None:
        if not is_distributed():
            raise RuntimeError("Distributed training must be initialized before creating ProcessGroupManager")
        sizes = {"tp": tp_size, "cp": cp_size, "pp": pp_size, "dp": dp_size, "ep": ep_size}
        for name, size in sizes.items():
            if size <= 0:
                raise ValueError(f"{name}_size must be positive, got {size}")
        self.global_rank = get_rank()
        self.world_size = get_world_size()
        self.local_rank = int(os.environ.get("LOCAL_RANK", self.global_rank % self.world_size))
        expected = tp_size * cp_size * pp_size * dp_size * ep_size
        if self.world_size != expected:
            raise ValueError(
                f"World size ({self.world_size}) != TP ({tp_size}) * CP ({cp_size}) * PP ({pp_size}) * "
                f"DP ({dp_size}) * EP ({ep_size}) = {expected}.")
        self.ep_replicate_data = bool(ep_replicate_data)
        self._sizes = sizes
        self.grid = torch.arange(self.world_size).view(dp_size, pp_size, cp_size, ep_size, tp_size)

        rem = self.global_rank
        self.tp_rank = rem % tp_size
        rem //= tp_size
        self.ep_rank = rem % ep_size
        rem //= ep_size
        self.cp_rank = rem % cp_size
        rem //= cp_size
        self.pp_rank = rem % pp_size
        self.dp_rank = rem // pp_size
        self._coords = {"dp": self.dp_rank, "pp": self.pp_rank, "cp": self.cp_rank, "ep": self.ep_rank,
                        "tp": self.tp_rank}

        self._all_groups: list[Any] = []
        self._create_process_groups()
        self._initialize_group_properties()

    # ------------------------------------------------------------------ group construction
    def _ranks_along(self, axes: tuple[str, ...], fixed: dict[str, int]) -> list[int]:
        """Ranks obtained by varying ``axes`` (in grid order) with the other coordinates fixed."""
        index = tuple(slice(None) if a in axes else fixed[a] for a in _AXES)
        return self.grid[index].reshape(-1).tolist()

    def _make_family(self, axes: tuple[str, ...]):
        """Creates one group per slice of the grid along ``axes`` (collective on every rank) and
        returns ``(my_group, my_ranks)``."""
        others = [a for a in _AXES if a not in axes]
        mine, mine_ranks = None, None
        for combo in product(*[range(self._sizes[a]) for a in others]):
            fixed = dict(zip(others, combo))
            ranks = self._ranks_along(axes, fixed)
            group = new_group(ranks=ranks)
            self._all_groups.append(group)
            if all(self._coords[a] == fixed[a] for a in others):
                mine, mine_ranks = group, ranks
        return mine, mine_ranks

    def _create_process_groups(self) -> None:
        self.tp_group, self.tp_group_ids = self._make_family(("tp",))
        self.cp_group, self.cp_group_ids = self._make_family(("cp",))
        self.pp_group, self.pp_group_ids = self._make_family(("pp",))
        self.ep_group, self.ep_group_ids = self._make_family(("ep",))
        self.dp_group, self.dp_group_ids = self._make_family(("dp",))
        self.cp_dp_group, self.cp_dp_group_ids = self._make_family(("dp", "cp"))
        self.pp_dp_group, self.pp_dp_group_ids = self._make_family(("dp", "pp"))
        if self._sizes["ep"] > 1:
            self.cp_dp_ep_group, self.cp_dp_ep_group_ids = self._make_family(("dp", "cp", "ep"))
        else:
            self.cp_dp_ep_group, self.cp_dp_ep_group_ids = self.cp_dp_group, self.cp_dp_group_ids
        # one model replica = everything except DP/CP (and EP when EP ranks hold distinct data,
        # because then each EP rank owns different experts of the same replica anyway)
        self.mp_group, self.mp_group_ids = self._make_family(("pp", "ep", "tp"))

    def _initialize_group_properties(self) -> None:
        s = self._sizes
        self.tp_world_size, self.cp_world_size, self.pp_world_size = s["tp"], s["cp"], s["pp"]
        self.dp_world_size, self.ep_world_size = s["dp"], s["ep"]
        for ax in ("tp", "cp", "pp", "dp", "ep"):
            ids = getattr(self, f"{ax}_group_ids")
            setattr(self, f"{ax}_first_rank", ids[0])
            setattr(self, f"{ax}_last_rank", ids[-1])
        # CP ring neighbours
        self.cp_send_rank = self.cp_group_ids[(self.cp_rank + 1) % self.cp_world_size]
        self.cp_recv_rank = self.cp_group_ids[(self.cp_rank - 1) % self.cp_world_size]
        # PP neighbours
        self.pp_is_first_stage = self.pp_rank == 0
        self.pp_is_last_stage = self.pp_rank == self.pp_world_size - 1
        self.pp_next_rank = None if self.pp_is_last_stage else self.pp_group_ids[self.pp_rank + 1]
        self.pp_prev_rank = None if self.pp_is_first_stage else self.pp_group_ids[self.pp_rank - 1]
        self.cp_dp_world_size = self.cp_world_size * self.dp_world_size
        self.pp_dp_world_size = self.pp_world_size * self.dp_world_size
        # gradient synchronisation domains (see module docstring)
        if self.ep_world_size > 1 and not self.ep_replicate_data:
            self.grad_sync_group, self.grad_sync_group_ids = self.cp_dp_ep_group, self.cp_dp_ep_group_ids
        else:
            self.grad_sync_group, self.grad_sync_group_ids = self.cp_dp_group, self.cp_dp_group_ids
        self.grad_sync_world_size = len(self.grad_sync_group_ids)
        self.expert_grad_sync_group, self.expert_grad_sync_group_ids = self.cp_dp_group, self.cp_dp_group_ids
        self.expert_grad_sync_world_size = self.cp_dp_world_size
        # data-loading replica coordinates
        if self.ep_world_size > 1 and not self.ep_replicate_data:
            self.data_rank = self.dp_rank * self.ep_world_size + self.ep_rank
            self.data_world_size = self.dp_world_size * self.ep_world_size
        else:
            self.data_rank, self.data_world_size = self.dp_rank, self.dp_world_size

    # ------------------------------------------------------------------ info
    def get_info(self) -> dict[str, Any]:
        return {
            "global_rank": self.global_rank, "world_size": self.world_size, "local_rank": self.local_rank,
            "tp_rank": self.tp_rank, "tp_world_size": self.tp_world_size,
            "cp_rank": self.cp_rank, "cp_world_size": self.cp_world_size,
            "pp_rank": self.pp_rank, "pp_world_size": self.pp_world_size,
            "dp_rank": self.dp_rank, "dp_world_size": self.dp_world_size,
            "ep_rank": self.ep_rank, "ep_world_size": self.ep_world_size,
            "pp_is_first_stage": self.pp_is_first_stage, "pp_is_last_stage": self.pp_is_last_stage,
            "grad_sync_world_size": self.grad_sync_world_size, "ep_replicate_data": self.ep_replicate_data,
        }

    def __str__(self) -> str:
        return (f"TP({self.tp_world_size})-CP({self.cp_world_size})-PP({self.pp_world_size})-"
                f"DP({self.dp_world_size})-EP({self.ep_world_size})-Rank({self.global_rank})")

    def __repr__(self) -> str:
        return (f"ProcessGroupManager(tp={self.tp_rank}/{self.tp_world_size}, cp={self.cp_rank}/{self.cp_world_size}, "
                f"pp={self.pp_rank}/{self.pp_world_size}, dp={self.dp_rank}/{self.dp_world_size}, "
                f"ep={self.ep_rank}/{self.ep_world_size}, global={self.global_rank}/{self.world_size})")

    def cleanup(self) -> None:
        """Destroys every group this manager created."""
        import torch.distributed as dist

        seen = set()
        for g in self._all_groups:
            if g is None or id(g) in seen:
                continue
            seen.add(id(g))
            try:
                dist.destroy_process_group(g)
            except Exception:
                pass
        self._all_groups.clear()


class ProcessGroupManagerProxy:
    """Falsy until :func:`setup_process_group_manager` installs an instance; ``if pgm:`` is the idiom for
    "are we distributed?" (reference ``process_group.py:359-405``)."""

    def __init__(self) -> None:
        object.__setattr__(self, "_instance", None)

    def __getattr__(self, name: str):
        inst = object.__getattribute__(self, "_instance")
        if inst is None:
            raise AttributeError(
                f"process group manager is not initialised (accessing '{name}'); call setup_process_group_manager()")
        return getattr(inst, name)

    def __setattr__(self, name: str, value) -> None:
        if name == "_instance":
            object.__setattr__(self, name, value)
        else:
            inst = object.__getattribute__(self, "_instance")
            if inst is None:
                raise AttributeError("process group manager is not initialised")
            setattr(inst, name, value)

    def __bool__(self) -> bool:
        return object.__getattribute__(self, "_instance") is not None

    def __repr__(self) -> str:
        inst = object.__getattribute__(self, "_instance")
        return "ProcessGroupManagerProxy(<uninitialised>)" if inst is None else repr(inst)

    def __str__(self) -> str:
        inst = object.__getattribute__(self, "_instance")
        return "ProcessGroupManagerProxy(<uninitialised>)" if inst is None else str(inst)


process_group_manager = ProcessGroupManagerProxy()


def setup_process_group_manager(tp_size: int, cp_size: int, pp_size: int, dp_size: int, ep_size: int = 1,
                                ep_replicate_data: bool = False) -> ProcessGroupManager:
    """Create the manager (collective: every rank must call it with the same sizes) and install it behind the global ``process_group_manager`` proxy."""
    inst = ProcessGroupManager(tp_size, cp_size, pp_size, dp_size, ep_size, ep_replicate_data)
    process_group_manager._instance = inst
    return inst


def get_process_group_manager() -> ProcessGroupManager | None:
    """The installed manager, or ``None`` before ``setup_process_group_manager`` / in a single-process run."""
    return object.__getattribute__(process_group_manager, "_instance")
