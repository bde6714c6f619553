"""Picklable-object collectives (reference: ``scaletorch/dist/object_ops.py:26-314``)."""
from __future__ import annotations

import pickle
from typing import Any

import torch
import torch.distributed as torch_dist
from torch import Tensor

from .utils import get_comm_device, get_default_group, get_rank, get_world_size


def _object_to_tensor(obj: Any) -> tuple[Tensor, Tensor]:
    data = pickle.dumps(obj)
    t = torch.frombuffer(bytearray(data), dtype=torch.uint8).clone()
    return t, torch.tensor([t.numel()], dtype=torch.long)


def _tensor_to_object(tensor: Tensor, tensor_size: int) -> Any:
    return pickle.loads(tensor.cpu().numpy().tobytes()[:tensor_size])


def broadcast_object_list(data: list[Any], src: int = 0, group=None) -> None:
    """Broadcasts the picklable objects of ``data`` from ``src`` in place."""
    if not isinstance(data, list):
        raise TypeError("data should be a list")
    if get_world_size(group) <= 1:
        return
    group = group or get_default_group()
    dev = get_comm_device(group)
    gsrc = torch_dist.get_global_rank(group, src)
    is_src = get_rank(group) == src
    if is_src:
        pairs = [_object_to_tensor(o) for o in data]
        sizes = torch.cat([s for _, s in pairs]).to(dev)
    else:
        sizes = torch.empty(len(data), dtype=torch.long, device=dev)
    torch_dist.broadcast(sizes, gsrc, group)
    if is_src:
        payload = torch.cat([t for t, _ in pairs]).to(dev)
    else:
        payload = torch.empty(int(sizes.sum().item()), dtype=torch.uint8, device=dev)
    torch_dist.broadcast(payload, gsrc, group)
    if not is_src:
        off = 0
        for i, n in enumerate(sizes.tolist()):
            data[i] = _tensor_to_object(payload[off:off + n], n)
            off += n


def all_gather_object(data: Any, group=None) -> list[Any]:
    """Every rank receives the list of all ranks' picklable ``data`` (sizes may differ per rank)."""
    world = get_world_size(group)
    if world <= 1:
        return [data]
    group = group or get_default_group()
    dev = get_comm_device(group)
    t, size = _object_to_tensor(data)
    sizes = torch.empty(world, dtype=torch.long, device=dev)
    torch_dist.all_gather_into_tensor(sizes, size.to(dev), group)
    mx = int(sizes.max().item())
    padded = torch.zeros(mx, dtype=torch.uint8, device=dev)
    padded[: t.numel()] = t.to(dev)
    out = torch.empty(world * mx, dtype=torch.uint8, device=dev)
    torch_dist.all_gather_into_tensor(out, padded, group)
    return [_tensor_to_object(out[i * mx:(i + 1) * mx], int(sizes[i].item())) for i in range(world)]


def gather_object(data: Any, dst: int = 0, group=None) -> list[Any] | None:
    """Group-local rank ``dst`` receives the list of all ranks' picklable ``data``; the others get ``None``."""
    world = get_world_size(group)
    if world <= 1:
        return [data]
    objs = all_gather_object(data, group)
    return objs if get_rank(group or get_default_group()) == dst else None
