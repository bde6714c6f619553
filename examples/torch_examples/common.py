"""Shared bring-up for the torch_examples (NCCL on GPUs, gloo on CPU) and a synthetic MNIST-shaped dataset (there is no
network access for downloads; pass ``--data-dir`` with torchvision's MNIST files to use the real one)."""
import os

import torch
import torch.distributed as dist
from torch.utils.data import Dataset


def setup():
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29513")
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if torch.cuda.is_available():
        dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
        torch.cuda.set_device(dev)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    else:
        dev = torch.device("cpu")
        dist.init_process_group("gloo", rank=rank, world_size=world)
    return rank, world, dev


class SyntheticDigits(Dataset):
    """28×28 images whose class is recoverable (a bright square whose position encodes the label) so accuracy is meaningful."""

    def __init__(self, n: int = 4096, seed: int = 0):
        g = torch.Generator().manual_seed(seed)
        self.y = torch.randint(0, 10, (n,), generator=g)
        self.x = 0.1 * torch.randn(n, 1, 28, 28, generator=g)
        for i, c in enumerate(self.y.tolist()):
            r, col = 2 + 5 * (c // 5) * 2, 2 + 5 * (c % 5)
            self.x[i, 0, r:r + 5, col:col + 5] += 1.0

    def __len__(self):
        return len(self.y)

    def __getitem__(self, i):
        return self.x[i], self.y[i]


def mnist(data_dir: str | None, train: bool):
    if data_dir:
        from torchvision import datasets, transforms

        return datasets.MNIST(data_dir, train=train, download=False,
                              transform=transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.1307,), (0.3081,))]))
    return SyntheticDigits(4096 if train else 1024, seed=0 if train else 1)
