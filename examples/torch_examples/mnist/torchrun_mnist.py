"""DDP on MNIST launched by ``torchrun`` (rank / world size / master address from the environment), elastic-restart
friendly: a snapshot per epoch lets a restarted worker group continue where it stopped (reference:
``examples/torch_examples/mnist/torchrun_mnist.py``).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 examples/torch_examples/mnist/torchrun_mnist.py --epochs 3 --snapshot /tmp/mnist_snap.pt
"""
import argparse
import os
import sys

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch.nn.parallel import DistributedDataParallel as DDP
from torch.optim.lr_scheduler import StepLR
from torch.utils.data import DataLoader
from torch.utils.data.distributed import DistributedSampler

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", ".."))
from common import mnist, setup  # noqa: E402
from scaletorch_b200.models.lenet import LeNet  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=3)
    ap.add_argument("--batch-size", type=int, default=64)
    ap.add_argument("--lr", type=float, default=0.05)
    ap.add_argument("--gamma", type=float, default=0.7)
    ap.add_argument("--snapshot", default="")
    ap.add_argument("--data-dir", default=None)
    a = ap.parse_args()
    rank, world, dev = setup()                     # reads RANK / WORLD_SIZE / LOCAL_RANK / MASTER_* set by torchrun
    torch.manual_seed(0)
    raw = LeNet().to(dev)
    opt = torch.optim.SGD(raw.parameters(), lr=a.lr, momentum=0.9)
    sched = StepLR(opt, step_size=1, gamma=a.gamma)
    start = 0
    if a.snapshot and os.path.exists(a.snapshot):
        snap = torch.load(a.snapshot, map_location=dev, weights_only=False)
        raw.load_state_dict(snap["model"])
        opt.load_state_dict(snap["opt"])
        sched.load_state_dict(snap["sched"])
        start = snap["epoch"]
        if rank == 0:
            print(f"restart: continuing at epoch {start}")
    model = DDP(raw, device_ids=[dev.index] if dev.type == "cuda" else None)
    train, test = mnist(a.data_dir, True), mnist(a.data_dir, False)
    sampler = DistributedSampler(train, world, rank, shuffle=True)
    loader = DataLoader(train, batch_size=a.batch_size, sampler=sampler, pin_memory=dev.type == "cuda")
    for epoch in range(start, a.epochs):
        sampler.set_epoch(epoch)
        model.train()
        for x, y in loader:
            opt.zero_grad()
            F.cross_entropy(model(x.to(dev)), y.to(dev)).backward()
            opt.step()
        sched.step()
        model.eval()
        stat = torch.zeros(3, device=dev)
        with torch.no_grad():
            for x, y in DataLoader(test, batch_size=512, sampler=DistributedSampler(test, world, rank, shuffle=False)):
                out = model(x.to(dev))
                stat[0] += (out.argmax(1) == y.to(dev)).sum()
                stat[1] += len(y)
                stat[2] += F.cross_entropy(out, y.to(dev), reduction="sum")
        dist.all_reduce(stat)
        if rank == 0:
            print(f"epoch {epoch + 1}: test loss {stat[2].item() / stat[1].item():.4f} accuracy {100.0 * stat[0].item() / stat[1].item():.1f} %", flush=True)
            if a.snapshot:
                tmp = a.snapshot + ".tmp"
                torch.save({"model": raw.state_dict(), "opt": opt.state_dict(), "sched": sched.state_dict(), "epoch": epoch + 1}, tmp)
                os.replace(tmp, a.snapshot)
        dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
