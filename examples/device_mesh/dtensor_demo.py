"""DTensor basics: distribute / redistribute / to_local / full_tensor, Shard vs Replicate vs Partial, and a sharded matmul
whose result placement is derived by the DTensor dispatcher.

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 4 examples/device_mesh/dtensor_demo.py
"""
import os
import sys

import torch
from torch.distributed.device_mesh import init_device_mesh
from torch.distributed.tensor import Partial, Replicate, Shard, distribute_tensor

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from common import finish, ordered_print, setup  # noqa: E402

rank, world, dev = setup()
mesh = init_device_mesh(dev.type, (world,))
torch.manual_seed(0)
big = torch.arange(world * 2 * 4, dtype=torch.float32, device=dev).view(world * 2, 4)

rows = distribute_tensor(big, mesh, [Shard(0)])
ordered_print(f"Shard(0): local {tuple(rows.to_local().shape)} first row {rows.to_local()[0].tolist()}")
cols = rows.redistribute(mesh, [Shard(1)]) if 4 % world == 0 else rows        # all-to-all under the hood
ordered_print(f"redistributed to {cols.placements}: local {tuple(cols.to_local().shape)}")
rep = rows.redistribute(mesh, [Replicate()])                                  # all-gather
assert torch.equal(rep.to_local(), big)

# column-parallel x row-parallel matmul: the product is Partial (pending sum) until it is redistributed
A = distribute_tensor(torch.randn(8, 4 * world, device=dev), mesh, [Shard(1)])
B = distribute_tensor(torch.randn(4 * world, 8, device=dev), mesh, [Shard(0)])
C = A @ B
ordered_print(f"A@B placements {C.placements} (Partial = all-reduce still pending)")
assert isinstance(C.placements[0], Partial)
full = C.full_tensor()
assert torch.allclose(full, A.full_tensor() @ B.full_tensor(), atol=1e-4)
ordered_print("full_tensor() matches the unsharded product")
finish()
