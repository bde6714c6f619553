"""2-D parallelism with upstream APIs: FSDP2 over the ``dp`` mesh dimension × DTensor tensor/sequence parallelism with
``loss_parallel`` over the ``tp`` dimension (topics of the reference's ``examples/FSDP2/fsdp2_tp_llama2_main.py:91-150,183-193``).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 4 examples/FSDP2/fsdp2_tp_main.py --tp 2 --steps 5
"""
from __future__ import annotations

import argparse
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch.distributed.device_mesh import init_device_mesh
from torch.distributed.fsdp import fully_shard
from torch.distributed.tensor import Replicate, Shard
from torch.distributed.tensor.parallel import (ColwiseParallel, PrepareModuleInput, RowwiseParallel, SequenceParallel,
                                               loss_parallel, parallelize_module)

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from model import ModelArgs, Transformer  # noqa: E402
from utils import describe_sharding, setup, synthetic_batch  # noqa: E402


def apply_tp(model: Transformer, tp_mesh) -> None:
    """Megatron layout + sequence parallelism: activations between blocks are sharded on the sequence dimension."""
    parallelize_module(model, tp_mesh, {
        "tok_embeddings": RowwiseParallel(input_layouts=Replicate(), output_layouts=Shard(1)),
        "norm": SequenceParallel(),
        "output": ColwiseParallel(input_layouts=Shard(1), output_layouts=Shard(-1), use_local_output=False),  # vocab-sharded logits
    })
    for blk in model.layers:
        parallelize_module(blk, tp_mesh, {
            "attention_norm": SequenceParallel(),
            "attention": PrepareModuleInput(input_layouts=(Shard(1), None, None), desired_input_layouts=(Replicate(), None, None)),
            "attention.wq": ColwiseParallel(), "attention.wk": ColwiseParallel(), "attention.wv": ColwiseParallel(),
            "attention.wo": RowwiseParallel(output_layouts=Shard(1)),
            "ffn_norm": SequenceParallel(),
            "feed_forward": PrepareModuleInput(input_layouts=(Shard(1),), desired_input_layouts=(Replicate(),)),
            "feed_forward.w1": ColwiseParallel(), "feed_forward.w3": ColwiseParallel(),
            "feed_forward.w2": RowwiseParallel(output_layouts=Shard(1)),
        })


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--tp", type=int, default=2)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--seq", type=int, default=256)
    a = ap.parse_args()
    rank, world, device = setup()
    assert world % a.tp == 0
    mesh = init_device_mesh(device.type, (world // a.tp, a.tp), mesh_dim_names=("dp", "tp"))
    torch.manual_seed(0)
    args = ModelArgs()
    model = Transformer(args).to(device)
    model.init_weights()
    apply_tp(model, mesh["tp"])
    for blk in model.layers:
        fully_shard(blk, mesh=mesh["dp"])
    fully_shard(model, mesh=mesh["dp"])
    if rank == 0:
        print(f"mesh {mesh}; parameters are DTensors sharded on (dp, tp):")
    describe_sharding(model, rank)
    opt = torch.optim.AdamW(model.parameters(), lr=3e-4, foreach=True)
    dp_rank = mesh["dp"].get_local_rank()
    for step in range(a.steps):
        x, y = synthetic_batch(a.batch, a.seq, args.vocab_size, device, seed=1000 * dp_rank + step)   # same data within a TP group
        with loss_parallel():                   # cross-entropy on vocab-sharded logits without gathering them
            loss = F.cross_entropy(model(x).flatten(0, 1), y.flatten())
            loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        if rank == 0:
            print(f"step {step:3d} loss {loss.item():.4f}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
