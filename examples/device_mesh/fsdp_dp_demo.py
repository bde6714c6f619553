"""Hybrid sharding (HSDP): shard inside a node ('shard' dim), replicate across nodes ('replicate' dim) via a 2-D mesh.

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 4 examples/device_mesh/fsdp_dp_demo.py
"""
import os
import sys

import torch
import torch.nn as nn
from torch.distributed.device_mesh import init_device_mesh
from torch.distributed.fsdp import fully_shard

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from common import finish, ordered_print, setup  # noqa: E402

rank, world, dev = setup()
assert world % 2 == 0
mesh = init_device_mesh(dev.type, (2, world // 2), mesh_dim_names=("replicate", "shard"))
torch.manual_seed(0)
model = nn.Sequential(nn.Linear(32, 128), nn.GELU(), nn.Linear(128, 32)).to(dev)
fully_shard(model, mesh=mesh)            # 2-D mesh => parameters sharded over 'shard', gradients also all-reduced over 'replicate'
loss = model(torch.randn(4, 32, device=dev)).sum()
loss.backward()
w = model[0].weight
ordered_print(f"placements {w.placements} local {tuple(w.to_local().shape)} grad local {tuple(w.grad.to_local().shape)}")
finish()
