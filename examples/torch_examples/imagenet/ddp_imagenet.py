"""ImageNet-style DDP training loop (ResNet-ish CNN, channels-last, bf16 autocast, top-1/top-5 meters reduced across
ranks) on synthetic 3×64×64 images; point ``--data`` at an ImageFolder tree to train on real data (reference:
``examples/torch_examples/imagenet``).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 examples/torch_examples/imagenet/ddp_imagenet.py --iters 10
"""
import argparse
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F
from torch.nn.parallel import DistributedDataParallel as DDP

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from common import setup  # noqa: E402


def conv_bn(i, o, s):
    return nn.Sequential(nn.Conv2d(i, o, 3, s, 1, bias=False), nn.BatchNorm2d(o), nn.ReLU(inplace=True))


class Res(nn.Module):
    def __init__(self, c):
        super().__init__()
        self.a, self.b = conv_bn(c, c, 1), nn.Sequential(nn.Conv2d(c, c, 3, 1, 1, bias=False), nn.BatchNorm2d(c))

    def forward(self, x):
        return F.relu(x + self.b(self.a(x)))


def small_resnet(classes=100):
    return nn.Sequential(conv_bn(3, 32, 2), Res(32), conv_bn(32, 64, 2), Res(64), conv_bn(64, 128, 2), Res(128),
                         nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(128, classes))


def topk(logits, y, ks=(1, 5)):
    pred = logits.topk(max(ks), 1).indices
    hit = pred == y[:, None]
    return [hit[:, :k].any(1).float().sum() for k in ks]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--batch-size", type=int, default=32)
    ap.add_argument("--data", default=None)
    a = ap.parse_args()
    rank, world, dev = setup()
    torch.manual_seed(0)
    net = small_resnet().to(dev).to(memory_format=torch.channels_last)
    net = nn.SyncBatchNorm.convert_sync_batchnorm(net) if dev.type == "cuda" else net
    model = DDP(net, device_ids=[dev.index] if dev.type == "cuda" else None)
    opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True)
    sched = torch.optim.lr_scheduler.OneCycleLR(opt, max_lr=0.2, total_steps=a.iters)
    loader = None
    if a.data:
        from torchvision import datasets, transforms

        ds = datasets.ImageFolder(a.data, transforms.Compose([transforms.RandomResizedCrop(64), transforms.ToTensor()]))
        sampler = torch.utils.data.distributed.DistributedSampler(ds, world, rank)
        loader = iter(torch.utils.data.DataLoader(ds, a.batch_size, sampler=sampler, num_workers=4, pin_memory=True))
    meters = torch.zeros(3, device=dev)
    for it in range(a.iters):
        if loader is not None:
            x, y = next(loader)
        else:
            g = torch.Generator().manual_seed(it * world + rank)
            y = torch.randint(0, 100, (a.batch_size,), generator=g)
            x = torch.randn(a.batch_size, 3, 64, 64, generator=g) + (y.float() / 50 - 1)[:, None, None, None]
        x, y = x.to(dev, memory_format=torch.channels_last, non_blocking=True), y.to(dev, non_blocking=True)
        with torch.autocast(dev.type, dtype=torch.bfloat16):
            logits = model(x)
            loss = F.cross_entropy(logits, y)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        sched.step()
        t1, t5 = topk(logits.float(), y)
        meters += torch.stack([t1, t5, torch.tensor(float(len(y)), device=dev)])
        if (it + 1) % 10 == 0:
            m = meters.clone()
            dist.all_reduce(m)
            if rank == 0:
                print(f"iter {it + 1} loss {loss.item():.3f} top1 {100 * m[0] / m[2]:.1f} % top5 {100 * m[1] / m[2]:.1f} %", flush=True)
            meters.zero_()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
