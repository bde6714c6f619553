"""2-D mesh: FSDP2 over 'dp' composed with tensor parallelism over 'tp' on a toy MLP stack (see examples/FSDP2 for the
Llama version).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 4 examples/device_mesh/fsdp_tp_demo.py
"""
import os
import sys

import torch
import torch.nn as nn
from torch.distributed.device_mesh import init_device_mesh
from torch.distributed.fsdp import fully_shard
from torch.distributed.tensor.parallel import ColwiseParallel, RowwiseParallel, parallelize_module

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from common import finish, ordered_print, setup  # noqa: E402

rank, world, dev = setup()
assert world % 2 == 0
mesh = init_device_mesh(dev.type, (world // 2, 2), mesh_dim_names=("dp", "tp"))
torch.manual_seed(0)
model = nn.Sequential(*[nn.Sequential(nn.Linear(64, 256, bias=False), nn.ReLU(), nn.Linear(256, 64, bias=False)) for _ in range(3)]).to(dev)
for blk in model:
    parallelize_module(blk, mesh["tp"], {"0": ColwiseParallel(), "2": RowwiseParallel()})
    fully_shard(blk, mesh=mesh["dp"])
fully_shard(model, mesh=mesh["dp"])
opt = torch.optim.SGD(model.parameters(), lr=0.1)
for step in range(3):
    torch.manual_seed(100 + mesh["dp"].get_local_rank() + 10 * step)     # same batch inside a tp group, different across dp
    loss = model(torch.randn(8, 64, device=dev)).pow(2).mean()
    loss.backward()
    opt.step()
    opt.zero_grad()
w = model[0][0].weight
ordered_print(f"weight {tuple(w.shape)} placements {w.placements} local {tuple(w.to_local().shape)} loss {loss.item():.4f}")
finish()
