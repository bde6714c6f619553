"""DeviceMesh API tour: N-D mesh creation, named sub-meshes, per-dimension groups/ranks, and how the framework's own
5-D grid (``ProcessGroupManager``: dp × pp × cp × ep × tp, TP fastest) maps onto the same rank layout.

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 4 examples/device_mesh/device_mesh_api.py
"""
import os
import sys

import torch
import torch.distributed as dist
from torch.distributed.device_mesh import init_device_mesh

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from common import finish, ordered_print, setup  # noqa: E402

rank, world, dev = setup()
assert world % 2 == 0, "run with an even number of processes"
mesh = init_device_mesh(dev.type, (world // 2, 2), mesh_dim_names=("dp", "tp"))
ordered_print(f"coordinate {mesh.get_coordinate()}  dp-rank {mesh['dp'].get_local_rank()}  tp-rank {mesh['tp'].get_local_rank()}  "
              f"tp peers {dist.get_process_group_ranks(mesh['tp'].get_group())}  dp peers {dist.get_process_group_ranks(mesh['dp'].get_group())}")
t = torch.ones(1, device=dev) * rank
dist.all_reduce(t, group=mesh["tp"].get_group())
ordered_print(f"sum of ranks inside my tp group = {int(t.item())}")

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))
from scaletorch_b200.parallel.process_group import setup_process_group_manager  # noqa: E402

pgm = setup_process_group_manager(tp_size=2, cp_size=1, pp_size=1, dp_size=world // 2, ep_size=1)
ordered_print(f"ProcessGroupManager: tp_rank {pgm.tp_rank} dp_rank {pgm.dp_rank}  tp group {pgm.tp_group_ids}  dp group {pgm.dp_group_ids}")
assert pgm.tp_rank == mesh["tp"].get_local_rank() and pgm.dp_rank == mesh["dp"].get_local_rank()
finish()
