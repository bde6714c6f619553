"""Collect per-rank result lists on rank 0 (reference: ``scaletorch/dist/gather_utils.py:24-211``)."""
from __future__ import annotations

import os
import pickle
import shutil
import tempfile

from .object_ops import all_gather_object, broadcast_object_list
from .utils import barrier, get_rank, get_world_size


def collect_results(results: list, size: int, device: str = "cpu", tmpdir: str | None = None) -> list | None:
    """Gather per-rank partial result lists on rank 0, interleaved back into dataset order and cut to ``size``; via a shared directory (``cpu``) or object collectives (``gpu``)."""
    if device not in ("cpu", "gpu", "cuda"):
        raise NotImplementedError(f"device must be 'cpu' or 'gpu', but got {device}")
    if device == "cpu":
        return collect_results_cpu(results, size, tmpdir)
    assert tmpdir is None, "tmpdir is only used with device='cpu'"
    return collect_results_gpu(results, size)


def _interleave(parts: list[list], size: int) -> list:
    ordered = []
    for items in zip(*[p + [None] * (max(map(len, parts)) - len(p)) for p in parts]):
        ordered.extend(items)
    ordered = [x for x in ordered if x is not None] if len(ordered) > size else ordered
    return ordered[:size]


def collect_results_cpu(result_part: list, size: int, tmpdir: str | None = None) -> list | None:
    """``collect_results`` through pickle files in a directory every rank can see (created and broadcast by rank 0 when not given)."""
    rank, world = get_rank(), get_world_size()
    if world == 1:
        return result_part[:size]
    if tmpdir is None:
        holder = [tempfile.mkdtemp(prefix="stb_collect_") if rank == 0 else None]
        broadcast_object_list(holder, 0)
        tmpdir = holder[0]
    else:
        os.makedirs(tmpdir, exist_ok=True)
    with open(os.path.join(tmpdir, f"part_{rank}.pkl"), "wb") as f:
        pickle.dump(result_part, f, protocol=2)
    barrier()
    if rank != 0:
        return None
    parts = []
    for i in range(world):
        path = os.path.join(tmpdir, f"part_{i}.pkl")
        if not os.path.exists(path):
            raise FileNotFoundError(f"{tmpdir} is not a shared directory for rank {i}")
        with open(path, "rb") as f:
            parts.append(pickle.load(f))
    shutil.rmtree(tmpdir, ignore_errors=True)
    return _interleave(parts, size)


def collect_results_gpu(result_part: list, size: int) -> list | None:
    """``collect_results`` through ``all_gather_object``; returns the ordered list on rank 0 and ``None`` elsewhere."""
    rank, world = get_rank(), get_world_size()
    if world == 1:
        return result_part[:size]
    parts = all_gather_object(result_part)
    if rank != 0:
        return None
    return _interleave(parts, size)
