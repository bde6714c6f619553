"""Sequence parallelism: norm / dropout regions run on a sequence shard; the TP region is entered by an all-gather and
left by a reduce-scatter (what ``scaletorch_b200`` fuses into AG→GEMM and GEMM→RS kernels).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 examples/device_mesh/sequence_parallel_demo.py
"""
import os
import sys

import torch
import torch.nn as nn
from torch.distributed.device_mesh import init_device_mesh
from torch.distributed.tensor import Shard, distribute_tensor
from torch.distributed.tensor.parallel import ColwiseParallel, RowwiseParallel, SequenceParallel, parallelize_module

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from common import finish, ordered_print, setup  # noqa: E402


class Block(nn.Module):
    def __init__(self, d=64, f=128):
        super().__init__()
        self.norm = nn.LayerNorm(d)
        self.up, self.down = nn.Linear(d, f, bias=False), nn.Linear(f, d, bias=False)

    def forward(self, x):
        return self.down(torch.relu(self.up(self.norm(x))))


rank, world, dev = setup()
mesh = init_device_mesh(dev.type, (world,))
torch.manual_seed(0)
ref = Block().to(dev)
sp = Block().to(dev)
sp.load_state_dict(ref.state_dict())
parallelize_module(sp, mesh, {"norm": SequenceParallel(), "up": ColwiseParallel(input_layouts=Shard(1)),
                              "down": RowwiseParallel(output_layouts=Shard(1), use_local_output=False)})
x = torch.randn(2, 8 * world, 64, device=dev)
y = sp(distribute_tensor(x, mesh, [Shard(1)]))
ordered_print(f"output placements {y.placements}, local shape {tuple(y.to_local().shape)}")
assert torch.allclose(y.full_tensor(), ref(x), atol=1e-5)
finish()
