"""Megatron-style tensor parallelism of an MLP with ``parallelize_module`` next to the framework's own fused layers.

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 examples/device_mesh/tensor_parallel_demo.py
"""
import os
import sys

import torch
import torch.nn as nn
from torch.distributed.device_mesh import init_device_mesh
from torch.distributed.tensor.parallel import ColwiseParallel, RowwiseParallel, parallelize_module

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from common import finish, ordered_print, setup  # noqa: E402


class MLP(nn.Module):
    def __init__(self, d=64, f=256):
        super().__init__()
        self.up, self.down = nn.Linear(d, f, bias=False), nn.Linear(f, d, bias=False)

    def forward(self, x):
        return self.down(torch.relu(self.up(x)))


rank, world, dev = setup()
mesh = init_device_mesh(dev.type, (world,))
torch.manual_seed(0)
ref = MLP().to(dev)
tp = MLP().to(dev)
tp.load_state_dict(ref.state_dict())
parallelize_module(tp, mesh, {"up": ColwiseParallel(), "down": RowwiseParallel()})   # one all-reduce in fwd, one in bwd
x = torch.randn(8, 64, device=dev)
y_ref, y_tp = ref(x), tp(x)
assert torch.allclose(y_ref, y_tp, atol=1e-5)
y_tp.sum().backward()
ordered_print(f"up.weight local shard {tuple(tp.up.weight.to_local().shape)}, down.weight local shard {tuple(tp.down.weight.to_local().shape)}; outputs match")
finish()
