"""Single-device MNIST baseline -- the program the DDP / FSDP variants next to it are derived from (reference:
``examples/torch_examples/mnist/basic_mnist.py``).

    python examples/torch_examples/mnist/basic_mnist.py --epochs 2 [--data-dir /path/to/MNIST] [--dry-run] [--save-model]
"""
import argparse
import os
import sys

import torch
import torch.nn.functional as F
from torch.optim.lr_scheduler import StepLR
from torch.utils.data import DataLoader

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", ".."))
from common import mnist  # noqa: E402
from scaletorch_b200.models.lenet import LeNet  # noqa: E402


def train_epoch(model, loader, opt, dev, epoch, log_interval, dry_run):
    model.train()
    for i, (x, y) in enumerate(loader):
        x, y = x.to(dev), y.to(dev)
        opt.zero_grad()
        loss = F.cross_entropy(model(x), y)
        loss.backward()
        opt.step()
        if i % log_interval == 0:
            print(f"epoch {epoch} [{i * len(x)}/{len(loader.dataset)}] loss {loss.item():.4f}")
            if dry_run:
                return


@torch.no_grad()
def evaluate(model, loader, dev) -> float:
    model.eval()
    correct = total = 0
    loss = 0.0
    for x, y in loader:
        x, y = x.to(dev), y.to(dev)
        out = model(x)
        loss += F.cross_entropy(out, y, reduction="sum").item()
        correct += (out.argmax(1) == y).sum().item()
        total += len(y)
    print(f"test: loss {loss / total:.4f} accuracy {correct}/{total} ({100.0 * correct / total:.1f} %)")
    return correct / total


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--batch-size", type=int, default=64)
    ap.add_argument("--test-batch-size", type=int, default=512)
    ap.add_argument("--lr", type=float, default=0.05)
    ap.add_argument("--gamma", type=float, default=0.7, help="learning-rate decay per epoch")
    ap.add_argument("--log-interval", type=int, default=20)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--dry-run", action="store_true")
    ap.add_argument("--save-model", action="store_true")
    ap.add_argument("--no-accel", action="store_true")
    ap.add_argument("--data-dir", default=None)
    a = ap.parse_args()
    torch.manual_seed(a.seed)
    dev = torch.device("cuda" if torch.cuda.is_available() and not a.no_accel else "cpu")
    train = DataLoader(mnist(a.data_dir, True), batch_size=a.batch_size, shuffle=True, pin_memory=dev.type == "cuda")
    test = DataLoader(mnist(a.data_dir, False), batch_size=a.test_batch_size)
    model = LeNet().to(dev)
    opt = torch.optim.SGD(model.parameters(), lr=a.lr, momentum=0.9)
    sched = StepLR(opt, step_size=1, gamma=a.gamma)
    for epoch in range(1, a.epochs + 1):
        train_epoch(model, train, opt, dev, epoch, a.log_interval, a.dry_run)
        evaluate(model, test, dev)
        sched.step()
    if a.save_model:
        torch.save(model.state_dict(), "mnist_lenet.pt")


if __name__ == "__main__":
    main()
