"""Tensor collectives with a single-process fast path (reference:
``scaletorch/dist/collective_ops.py:33-960``).

Every function is a no-op (or identity) when ``world_size == 1``; tensors are moved to the backend's
communication device and back when they live elsewhere (gloo groups + CUDA tensors).  Unlike the
reference the flat ``*_into_tensor`` / ``reduce_scatter_tensor`` c10d variants are used where the
layout allows, avoiding the list + ``torch.cat`` copies (``collective_ops.py:603-605``).
"""
from __future__ import annotations

from types import GeneratorType
from typing import Any

import numpy as np
import torch
import torch.distributed as torch_dist
from torch import Tensor
from torch.distributed import ProcessGroup

from ._reduce_op import _get_reduce_op
from .utils import (cast_data_device, get_backend, get_comm_device, get_data_device, get_default_group,
                    get_rank, get_world_size)


def _flatten_dense_tensors(tensors: list[Tensor]) -> Tensor:
    return torch.cat([t.reshape(-1) for t in tensors]) if len(tensors) > 0 else torch.empty(0)


def _unflatten_dense_tensors(flat: Tensor, tensors: list[Tensor]) -> list[Tensor]:
    outs, off = [], 0
    for t in tensors:
        n = t.numel()
        outs.append(flat[off:off + n].view_as(t))
        off += n
    return outs


def _take_tensors(tensors, size_limit: int):
    """Yields lists of same-dtype tensors whose total byte size stays under ``size_limit``."""
    buckets: dict[Any, tuple[list[Tensor], int]] = {}
    for t in tensors:
        key = t.dtype
        lst, size = buckets.get(key, ([], 0))
        nbytes = t.numel() * t.element_size()
        if lst and size + nbytes > size_limit:
            yield lst
            lst, size = [], 0
        lst.append(t)
        buckets[key] = (lst, size + nbytes)
    for lst, _ in buckets.values():
        if lst:
            yield lst


def _group(group):
    return group if group is not None else get_default_group()


def all_reduce(data: Tensor, op: str = "sum", group: ProcessGroup | None = None, async_op: bool = False):
    """In-place all-reduce.  ``'mean'`` = sum then divide.  Returns the work handle if ``async_op``."""
    world = get_world_size(group)
    if world <= 1:
        return None
    group = _group(group)
    in_dev = get_data_device(data)
    comm_dev = get_comm_device(group)
    moved = cast_data_device(data, comm_dev)
    if op.lower() == "mean":
        if async_op and moved is data:
            data.div_(world)  # pre-divide so the async result needs no post-processing
            return torch_dist.all_reduce(data, _get_reduce_op("sum"), group, async_op=True)
        torch_dist.all_reduce(moved, _get_reduce_op("sum"), group)
        moved = moved / world if not moved.is_floating_point() else moved.div_(world)
    else:
        if async_op and moved is data:
            return torch_dist.all_reduce(data, _get_reduce_op(op), group, async_op=True)
        torch_dist.all_reduce(moved, _get_reduce_op(op), group)
    if moved is not data:
        cast_data_device(moved, in_dev, out=data)
    return None


def all_gather(data: Tensor, group: ProcessGroup | None = None) -> list[Tensor]:
    """Every rank receives the list of all ranks' ``data`` (equal shapes), in rank order, on the input's device."""
    world = get_world_size(group)
    if world <= 1:
        return [data]
    group = _group(group)
    in_dev = get_data_device(data)
    comm_dev = get_comm_device(group)
    moved = cast_data_device(data, comm_dev).contiguous()
    flat = torch.empty((world,) + tuple(moved.shape), dtype=moved.dtype, device=comm_dev)
    torch_dist.all_gather_into_tensor(flat.view(-1), moved.view(-1), group)
    return [cast_data_device(flat[i], in_dev) for i in range(world)]


def gather(data: Tensor, dst: int = 0, group: ProcessGroup | None = None) -> list[Tensor | None]:
    """Rank ``dst`` (group-local) receives the list of all ranks' ``data``; the others get ``[]``.  NCCL has no gather, so it is an all-gather there."""
    world = get_world_size(group)
    if world <= 1:
        return [data]
    group = _group(group)
    in_dev = get_data_device(data)
    comm_dev = get_comm_device(group)
    moved = cast_data_device(data, comm_dev)
    if get_backend(group) == "nccl":  # NCCL has no gather: emulate with all_gather
        out = all_gather(moved, group)
        return [cast_data_device(t, in_dev) for t in out] if get_rank(group) == dst else []
    is_dst = get_rank(group) == dst
    lst = [torch.empty_like(moved) for _ in range(world)] if is_dst else None
    torch_dist.gather(moved, lst, torch_dist.get_global_rank(group, dst), group)
    return [cast_data_device(t, in_dev) for t in lst] if is_dst else []


def broadcast(data: Tensor, src: int = 0, group: ProcessGroup | None = None) -> None:
    """Overwrite ``data`` on every rank with the value held by group-local rank ``src`` (in place)."""
    if get_world_size(group) <= 1:
        return
    group = _group(group)
    in_dev = get_data_device(data)
    comm_dev = get_comm_device(group)
    moved = cast_data_device(data, comm_dev).contiguous()
    torch_dist.broadcast(moved, torch_dist.get_global_rank(group, src), group)
    if moved is not data and get_rank(group) != src:
        cast_data_device(moved, in_dev, out=data)


def scatter(tensor_out: Tensor, scatter_list: list[Tensor] | None = None, src: int = 0,
            group: ProcessGroup | None = None) -> None:
    """Group-local rank ``src`` sends ``scatter_list[i]`` to rank ``i``; every rank receives into ``tensor_out``."""
    world = get_world_size(group)
    if world <= 1:
        if scatter_list:
            tensor_out.copy_(scatter_list[0])
        return
    group = _group(group)
    in_dev = get_data_device(tensor_out)
    comm_dev = get_comm_device(group)
    out = cast_data_device(tensor_out, comm_dev)
    is_src = get_rank(group) == src
    if is_src:
        if scatter_list is None or len(scatter_list) != world:
            raise ValueError("scatter_list must hold world_size tensors on the source rank")
        lst = [cast_data_device(t, comm_dev).contiguous() for t in scatter_list]
    else:
        lst = None
    gsrc = torch_dist.get_global_rank(group, src)
    if get_backend(group) == "nccl":
        torch_dist.scatter(out, lst, gsrc, group)
    else:
        torch_dist.scatter(out, lst, gsrc, group)
    if out is not tensor_out:
        cast_data_device(out, in_dev, out=tensor_out)


def reduce(data: Tensor, dst: int = 0, op: str = "sum", group: ProcessGroup | None = None) -> None:
    """Reduce ``data`` over the group into rank ``dst`` (in place there; other ranks' buffers are unspecified).  ``op`` = sum / mean / max / min / product."""
    world = get_world_size(group)
    if world <= 1:
        return
    group = _group(group)
    in_dev = get_data_device(data)
    comm_dev = get_comm_device(group)
    moved = cast_data_device(data, comm_dev)
    torch_dist.reduce(moved, torch_dist.get_global_rank(group, dst), _get_reduce_op(op), group)
    if get_rank(group) == dst:
        if op.lower() == "mean":
            moved = moved.div_(world) if moved.is_floating_point() else moved / world
        if moved is not data:
            cast_data_device(moved, in_dev, out=data)
        elif op.lower() == "mean" and not data.is_floating_point():
            data.copy_(moved)


def reduce_scatter(tensor_out: Tensor, input_list: list[Tensor] | None = None, op: str = "sum",
                   group: ProcessGroup | None = None) -> None:
    """Rank ``i`` receives the reduction over ranks of ``input_list[i]`` in ``tensor_out`` (all-reduce + slice on gloo)."""
    world = get_world_size(group)
    if world <= 1:
        if input_list:
            tensor_out.copy_(input_list[0])
        return
    group = _group(group)
    if input_list is None or len(input_list) != world:
        raise ValueError("input_list must hold world_size tensors")
    in_dev = get_data_device(tensor_out)
    comm_dev = get_comm_device(group)
    out = cast_data_device(tensor_out, comm_dev)
    flat_in = torch.stack([cast_data_device(t, comm_dev) for t in input_list]).contiguous()
    if get_backend(group) == "gloo":  # gloo lacks reduce_scatter: all-reduce + slice
        torch_dist.all_reduce(flat_in, _get_reduce_op(op), group)
        out.copy_(flat_in[get_rank(group)])
    else:
        torch_dist.reduce_scatter_tensor(out.contiguous().view(-1), flat_in.view(-1), _get_reduce_op(op), group)
    if op.lower() == "mean":
        out.div_(world)
    if out is not tensor_out:
        cast_data_device(out, in_dev, out=tensor_out)


def all_to_all(output_tensor_list: list[Tensor], input_tensor_list: list[Tensor],
               group: ProcessGroup | None = None) -> list[Tensor]:
    """Rank ``r`` sends ``input_tensor_list[p]`` to peer ``p`` and receives ``output_tensor_list[p]`` from it (p2p emulation on gloo)."""
    world = get_world_size(group)
    if world <= 1:
        for o, i in zip(output_tensor_list, input_tensor_list):
            o.copy_(i)
        return output_tensor_list
    group = _group(group)
    if len(output_tensor_list) != world or len(input_tensor_list) != world:
        raise ValueError("tensor lists must hold world_size tensors")
    in_dev = get_data_device(output_tensor_list)
    comm_dev = get_comm_device(group)
    outs = [cast_data_device(t, comm_dev) for t in output_tensor_list]
    ins = [cast_data_device(t, comm_dev).contiguous() for t in input_tensor_list]
    if get_backend(group) == "gloo":  # emulate with p2p
        rank = get_rank(group)
        reqs = []
        for peer in range(world):
            if peer == rank:
                outs[peer].copy_(ins[peer])
                continue
            g = torch_dist.get_global_rank(group, peer)
            reqs.append(torch_dist.isend(ins[peer], g, group=group))
            reqs.append(torch_dist.irecv(outs[peer], g, group=group))
        for r in reqs:
            r.wait()
    else:
        torch_dist.all_to_all(outs, ins, group)
    for o, t in zip(outs, output_tensor_list):
        if o is not t:
            cast_data_device(o, in_dev, out=t)
    return output_tensor_list


def sync_random_seed(group: ProcessGroup | None = None) -> int:
    """A random seed that is identical on every rank of ``group`` (rank 0 draws it)."""
    seed = int(np.random.randint(2**31))
    if get_world_size(group) <= 1:
        return seed
    group = _group(group)
    t = torch.tensor([seed if get_rank(group) == 0 else 0], dtype=torch.int64, device=get_comm_device(group))
    torch_dist.broadcast(t, torch_dist.get_global_rank(group, 0), group)
    return int(t.item())


def all_reduce_dict(data: dict[str, Tensor], op: str = "sum", group: ProcessGroup | None = None) -> None:
    """All-reduces every value of a dict with one flattened collective (keys must match across ranks)."""
    if not isinstance(data, dict):
        raise TypeError("data should be a dict")
    if get_world_size(group) <= 1 or not data:
        return
    keys = sorted(data.keys())
    tensors = [data[k] for k in keys]
    dtype = tensors[0].dtype
    flat = _flatten_dense_tensors([t.to(dtype) for t in tensors])
    all_reduce(flat, op=op, group=group)
    for k, piece in zip(keys, _unflatten_dense_tensors(flat, tensors)):
        data[k].copy_(piece.to(data[k].dtype))


def _all_reduce_coalesced(tensors: list[Tensor], bucket_size_mb: int = -1, op: str = "sum",
                          group: ProcessGroup | None = None) -> None:
    if bucket_size_mb > 0:
        buckets = list(_take_tensors(tensors, bucket_size_mb * 1024 * 1024))
    else:
        by_type: dict[Any, list[Tensor]] = {}
        for t in tensors:
            by_type.setdefault(t.dtype, []).append(t)
        buckets = list(by_type.values())
    for bucket in buckets:
        flat = _flatten_dense_tensors(bucket)
        all_reduce(flat, op=op, group=group)
        for t, synced in zip(bucket, _unflatten_dense_tensors(flat, bucket)):
            t.copy_(synced)


def all_reduce_params(params: list | GeneratorType, coalesce: bool = True, bucket_size_mb: int = -1,
                      op: str = "sum", group: ProcessGroup | None = None) -> None:
    """All-reduce a collection of parameters in place, optionally coalesced into flat buckets of ``bucket_size_mb``."""
    if get_world_size(group) <= 1:
        return
    tensors = [p.data for p in params]
    if coalesce:
        _all_reduce_coalesced(tensors, bucket_size_mb, op=op, group=group)
    else:
        for t in tensors:
            all_reduce(t, op=op, group=group)
