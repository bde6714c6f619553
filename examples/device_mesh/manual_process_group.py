"""The same 2-D layout built by hand with ``dist.new_group`` -- what DeviceMesh (and ``ProcessGroupManager``) automate.
Every rank must create every group, in the same order, even the ones it is not a member of.

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 4 examples/device_mesh/manual_process_group.py
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from common import finish, ordered_print, setup  # noqa: E402

rank, world, dev = setup()
assert world % 2 == 0
tp = 2
grid = torch.arange(world).view(world // tp, tp)
tp_group = dp_group = None
for row in grid.tolist():                   # tp groups: consecutive ranks (fastest-varying dimension -> same NVLink island)
    g = dist.new_group(row)
    if rank in row:
        tp_group = g
for col in grid.t().tolist():               # dp groups: strided ranks
    g = dist.new_group(col)
    if rank in col:
        dp_group = g
t = torch.tensor([float(rank)], device=dev)
a, b = t.clone(), t.clone()
dist.all_reduce(a, group=tp_group)
dist.all_reduce(b, group=dp_group)
ordered_print(f"tp peers {dist.get_process_group_ranks(tp_group)} sum {int(a.item())} | dp peers {dist.get_process_group_ranks(dp_group)} sum {int(b.item())}")
finish()
