"""DDP on MNIST launched with ``mp.spawn`` from ONE python process (no torchrun): the parent picks a free port, spawns one
worker per GPU (or ``--nprocs`` CPU workers over gloo), each worker builds its process group from its spawn index
(reference: ``examples/torch_examples/mnist/multigpu_mnist.py``).

    python examples/torch_examples/mnist/multigpu_mnist.py --epochs 2 [--nprocs 2]
"""
import argparse
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn.functional as F
from torch.nn.parallel import DistributedDataParallel as DDP
from torch.utils.data import DataLoader
from torch.utils.data.distributed import DistributedSampler

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", ".."))
from common import mnist  # noqa: E402
from scaletorch_b200.models.lenet import LeNet  # noqa: E402


def worker(rank: int, world: int, port: int, a) -> None:
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    use_cuda = torch.cuda.is_available() and torch.cuda.device_count() >= world
    dev = torch.device("cuda", rank) if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(dev)
    dist.init_process_group("nccl" if use_cuda else "gloo", rank=rank, world_size=world)
    torch.manual_seed(a.seed)                      # identical initial weights; DDP broadcasts rank 0's anyway
    model = DDP(LeNet().to(dev), device_ids=[rank] if use_cuda else None)
    train, test = mnist(a.data_dir, True), mnist(a.data_dir, False)
    sampler = DistributedSampler(train, world, rank, shuffle=True, seed=a.seed)
    loader = DataLoader(train, batch_size=a.batch_size, sampler=sampler)
    opt = torch.optim.SGD(model.parameters(), lr=a.lr, momentum=0.9)
    for epoch in range(a.epochs):
        sampler.set_epoch(epoch)
        model.train()
        for x, y in loader:
            opt.zero_grad()
            F.cross_entropy(model(x.to(dev)), y.to(dev)).backward()
            opt.step()
        model.eval()
        stat = torch.zeros(2, device=dev)
        with torch.no_grad():
            for x, y in DataLoader(test, batch_size=512, sampler=DistributedSampler(test, world, rank, shuffle=False)):
                stat[0] += (model(x.to(dev)).argmax(1) == y.to(dev)).sum()
                stat[1] += len(y)
        dist.all_reduce(stat)
        if rank == 0:
            print(f"epoch {epoch + 1}: accuracy {100.0 * stat[0].item() / stat[1].item():.1f} % over {int(stat[1].item())} samples", flush=True)
    if rank == 0 and a.save_model:
        torch.save(model.module.state_dict(), "mnist_lenet_ddp.pt")
    dist.barrier()
    dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--batch-size", type=int, default=64)
    ap.add_argument("--lr", type=float, default=0.05)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--nprocs", type=int, default=0, help="0 = one per visible GPU (2 CPU workers without GPUs)")
    ap.add_argument("--save-model", action="store_true")
    ap.add_argument("--data-dir", default=None)
    a = ap.parse_args()
    world = a.nprocs or (torch.cuda.device_count() if torch.cuda.is_available() else 2)
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(worker, args=(world, port, a), nprocs=world, join=True)


if __name__ == "__main__":
    main()
