"""Llama-2 on a 2-D device mesh: tensor parallelism (+ sequence parallelism around the norms) inside a node, FSDP2 across
the data-parallel dimension (reference: ``examples/FSDP2/fsdp2_tp_llama2_main.py``).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 4 examples/FSDP2/fsdp2_tp_llama2_main.py --tp 2 --size tiny --steps 10
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
                                               parallelize_module)

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from data import build_loader  # noqa: E402
from llama2 import Llama2, Llama2Config  # noqa: E402
from utils import describe_sharding, setup  # noqa: E402


def tp_plan_for_block():
    """Megatron-style plan: q/k/v and gate/up column-wise, o/down row-wise; activations between blocks stay sharded along
    the sequence (``Shard(1)``), are gathered in front of attention / MLP and reduce-scattered behind them."""
    return {
        "input_layernorm": SequenceParallel(),
        "self_attn": PrepareModuleInput(input_layouts=(Shard(1), None, None), desired_input_layouts=(Replicate(), None, None)),
        "self_attn.q_proj": ColwiseParallel(),
        "self_attn.k_proj": ColwiseParallel(),
        "self_attn.v_proj": ColwiseParallel(),
        "self_attn.o_proj": RowwiseParallel(output_layouts=Shard(1)),
        "post_attention_layernorm": SequenceParallel(),
        "mlp": PrepareModuleInput(input_layouts=(Shard(1),), desired_input_layouts=(Replicate(),)),
        "mlp.gate_proj": ColwiseParallel(),
        "mlp.up_proj": ColwiseParallel(),
        "mlp.down_proj": RowwiseParallel(output_layouts=Shard(1)),
    }


def parallelize(model: Llama2, mesh):
    tp_mesh, dp_mesh = mesh["tp"], mesh["dp"]
    if tp_mesh.size() > 1:
        parallelize_module(model, tp_mesh, {
            "embed_tokens": RowwiseParallel(input_layouts=Replicate(), output_layouts=Shard(1)),
            "norm": SequenceParallel(),
            "lm_head": ColwiseParallel(input_layouts=Shard(1), output_layouts=Replicate()),
        })
        for blk in model.layers:
            parallelize_module(blk, tp_mesh, tp_plan_for_block())
    if dp_mesh.size() > 1:
        for blk in model.layers:
            fully_shard(blk, mesh=dp_mesh)
        fully_shard(model, mesh=dp_mesh)
    return model


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--tp", type=int, default=2)
    ap.add_argument("--size", default="tiny")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--seq", type=int, default=128)
    ap.add_argument("--lr", type=float, default=3e-4)
    a = ap.parse_args()
    rank, world, device = setup()
    if world % a.tp != 0:
        raise SystemExit(f"world size {world} is not a multiple of --tp {a.tp}")
    mesh = init_device_mesh(device.type, (world // a.tp, a.tp), mesh_dim_names=("dp", "tp"))
    cfg = Llama2Config.named(a.size)
    if cfg.n_kv_heads % a.tp != 0:
        raise SystemExit(f"--tp {a.tp} must divide the {cfg.n_kv_heads} KV heads")
    torch.manual_seed(0)                                  # same initial weights on every rank before they are distributed
    model = Llama2(cfg).to(device)
    parallelize(model, mesh)
    opt = torch.optim.AdamW(model.parameters(), lr=a.lr, foreach=True)
    # ranks of one TP group must see the SAME batch: the data rank is the dp coordinate
    loader = build_loader("synthetic", None, cfg.vocab_size, a.batch, a.seq, mesh["dp"].get_local_rank(), mesh["dp"].size())
    if rank == 0:
        print(f"mesh dp={mesh['dp'].size()} x tp={mesh['tp'].size()} on {device.type}")
    describe_sharding(model, rank)
    for step in range(a.steps):
        x, y = next(loader)
        logits = model(x.to(device))
        loss = F.cross_entropy(logits.float().flatten(0, 1), y.to(device).flatten())
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        if rank == 0:
            print(f"step {step:3d} loss {loss.item():.4f}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
