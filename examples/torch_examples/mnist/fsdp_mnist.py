"""The same MNIST job with FSDP1 (``FullyShardedDataParallel``, size-based auto-wrap) and a rank-0 full state-dict save.

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 examples/torch_examples/mnist/fsdp_mnist.py --epochs 1
"""
import argparse
import functools
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch.distributed.fsdp import FullStateDictConfig, StateDictType
from torch.distributed.fsdp import FullyShardedDataParallel as FSDP
from torch.distributed.fsdp.wrap import size_based_auto_wrap_policy
from torch.utils.data import DataLoader
from torch.utils.data.distributed import DistributedSampler

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", ".."))
from common import mnist, setup  # noqa: E402
from scaletorch_b200.models.lenet import LeNet  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=1)
    ap.add_argument("--batch-size", type=int, default=64)
    ap.add_argument("--save", default="")
    ap.add_argument("--data-dir", default=None)
    a = ap.parse_args()
    rank, world, dev = setup()
    torch.manual_seed(0)
    policy = functools.partial(size_based_auto_wrap_policy, min_num_params=20_000)
    fsdp1 = dev.type == "cuda"               # FSDP1 requires an accelerator; on CPU the same job runs with FSDP2
    if fsdp1:
        model = FSDP(LeNet().to(dev), auto_wrap_policy=policy, device_id=dev)
    else:
        from torch.distributed.fsdp import fully_shard

        model = LeNet()
        for m in (model.fc1, model.conv2):
            fully_shard(m)
        fully_shard(model)
    train = mnist(a.data_dir, True)
    sampler = DistributedSampler(train, world, rank, shuffle=True)
    opt = torch.optim.Adadelta(model.parameters(), lr=1.0)
    for epoch in range(a.epochs):
        sampler.set_epoch(epoch)
        tot = torch.zeros(2, device=dev)
        for x, y in DataLoader(train, batch_size=a.batch_size, sampler=sampler):
            loss = F.cross_entropy(model(x.to(dev)), y.to(dev))
            opt.zero_grad()
            loss.backward()
            opt.step()
            tot += torch.tensor([loss.item() * len(y), len(y)], device=dev)
        dist.all_reduce(tot)
        if rank == 0:
            print(f"epoch {epoch} mean loss {tot[0].item() / tot[1].item():.4f}", flush=True)
    if a.save:
        if fsdp1:
            with FSDP.state_dict_type(model, StateDictType.FULL_STATE_DICT, FullStateDictConfig(offload_to_cpu=True, rank0_only=True)):
                sd = model.state_dict()
        else:
            from torch.distributed.checkpoint.state_dict import StateDictOptions, get_model_state_dict

            sd = get_model_state_dict(model, options=StateDictOptions(full_state_dict=True, cpu_offload=True))
        if rank == 0:
            torch.save(sd, a.save)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
