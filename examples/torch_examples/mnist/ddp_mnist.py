"""DistributedDataParallel on MNIST-shaped data with the package's LeNet (reference: ``examples/torch_examples/mnist``).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 examples/torch_examples/mnist/ddp_mnist.py --epochs 2
"""
import argparse
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch.nn.parallel import DistributedDataParallel as DDP
from torch.utils.data import DataLoader
from torch.utils.data.distributed import DistributedSampler

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", ".."))
from common import mnist, setup  # noqa: E402
from scaletorch_b200.models.lenet import LeNet  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--batch-size", type=int, default=64)
    ap.add_argument("--lr", type=float, default=0.05)
    ap.add_argument("--data-dir", default=None)
    a = ap.parse_args()
    rank, world, dev = setup()
    torch.manual_seed(0)
    model = DDP(LeNet().to(dev), device_ids=[dev.index] if dev.type == "cuda" else None)
    train, test = mnist(a.data_dir, True), mnist(a.data_dir, False)
    sampler = DistributedSampler(train, num_replicas=world, rank=rank, shuffle=True)
    loader = DataLoader(train, batch_size=a.batch_size, sampler=sampler)
    opt = torch.optim.SGD(model.parameters(), lr=a.lr, momentum=0.9)
    for epoch in range(a.epochs):
        sampler.set_epoch(epoch)                     # a different shuffle each epoch, consistent across ranks
        model.train()
        for x, y in loader:
            loss = F.cross_entropy(model(x.to(dev)), y.to(dev))
            opt.zero_grad()
            loss.backward()                          # DDP all-reduces gradient buckets while backward still runs
            opt.step()
        model.eval()
        stats = torch.zeros(2, device=dev)
        with torch.no_grad():
            for x, y in DataLoader(test, batch_size=256, sampler=DistributedSampler(test, world, rank, shuffle=False)):
                stats[0] += (model(x.to(dev)).argmax(1) == y.to(dev)).sum()
                stats[1] += len(y)
        dist.all_reduce(stats)
        if rank == 0:
            print(f"epoch {epoch} last-batch loss {loss.item():.4f} test accuracy {100 * stats[0].item() / stats[1].item():.1f} %", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
