"""DDP trainer with snapshot save / resume for the miniGPT example (reference:
``examples/torch_examples/minigpt/trainer.py``): DistributedSampler epochs, optional bf16 autocast, gradient clipping, a
snapshot (model, optimiser, finished epochs) written by rank 0 every ``save_every`` epochs and picked up by a relaunched
job (torchrun restart after a failure)."""
from __future__ import annotations

import os
from dataclasses import dataclass

import torch
import torch.distributed as dist
from torch.nn.parallel import DistributedDataParallel as DDP
from torch.utils.data import DataLoader
from torch.utils.data.distributed import DistributedSampler


@dataclass
class TrainerConfig:
    max_epochs: int = 2
    batch_size: int = 32
    data_loader_workers: int = 0
    grad_norm_clip: float = 1.0
    snapshot_path: str = "/tmp/minigpt_snapshot.pt"
    save_every: int = 1
    use_amp: bool = False
    max_steps_per_epoch: int | None = None      # cap for smoke runs


class Trainer:
    def __init__(self, cfg: TrainerConfig, model, optimizer, train_set, test_set, device, rank: int, world: int):
        self.cfg, self.device, self.rank, self.world = cfg, device, rank, world
        self.train_sampler = DistributedSampler(train_set, world, rank, shuffle=True)
        self.train_loader = DataLoader(train_set, batch_size=cfg.batch_size, sampler=self.train_sampler,
                                       num_workers=cfg.data_loader_workers, pin_memory=device.type == "cuda", drop_last=True)
        self.test_loader = DataLoader(test_set, batch_size=cfg.batch_size, num_workers=cfg.data_loader_workers,
                                      sampler=DistributedSampler(test_set, world, rank, shuffle=False)) if test_set else None
        self.optimizer = optimizer
        self.epochs_run = 0
        model = model.to(device)
        if os.path.exists(cfg.snapshot_path):
            self._load(model)
        self.model = DDP(model, device_ids=[device.index] if device.type == "cuda" else None)

    # ------------------------------------------------------------------ snapshots
    def _load(self, raw_model) -> None:
        snap = torch.load(self.cfg.snapshot_path, map_location=self.device, weights_only=False)
        raw_model.load_state_dict(snap["model"])
        self.optimizer.load_state_dict(snap["optimizer"])
        self.epochs_run = int(snap["epochs_run"])
        if self.rank == 0:
            print(f"resuming from snapshot at epoch {self.epochs_run}")

    def _save(self, epoch: int) -> None:
        tmp = self.cfg.snapshot_path + ".tmp"
        torch.save({"model": self.model.module.state_dict(), "optimizer": self.optimizer.state_dict(),
                    "epochs_run": epoch}, tmp)
        os.replace(tmp, self.cfg.snapshot_path)          # atomic: a crash never leaves a half-written snapshot
        print(f"snapshot saved at epoch {epoch}: {self.cfg.snapshot_path}")

    # ------------------------------------------------------------------ loops
    def _run_batch(self, x, y, train: bool) -> float:
        ctx = torch.autocast(self.device.type, dtype=torch.bfloat16) if self.cfg.use_amp else torch.autocast(self.device.type, enabled=False)
        with torch.set_grad_enabled(train), ctx:
            _, loss = self.model(x, y)
        if train:
            self.optimizer.zero_grad(set_to_none=True)
            loss.backward()
            torch.nn.utils.clip_grad_norm_(self.model.parameters(), self.cfg.grad_norm_clip)
            self.optimizer.step()
        return float(loss.detach())

    def _run_epoch(self, epoch: int, loader, train: bool) -> float:
        if train:
            self.train_sampler.set_epoch(epoch)
        self.model.train(train)
        total, n = 0.0, 0
        for i, (x, y) in enumerate(loader):
            if self.cfg.max_steps_per_epoch and i >= self.cfg.max_steps_per_epoch:
                break
            total += self._run_batch(x.to(self.device, non_blocking=True), y.to(self.device, non_blocking=True), train)
            n += 1
        t = torch.tensor([total, n], dtype=torch.float64, device=self.device)
        dist.all_reduce(t)
        return float(t[0] / t[1].clamp(min=1))

    def train(self) -> None:
        for epoch in range(self.epochs_run, self.cfg.max_epochs):
            tr = self._run_epoch(epoch, self.train_loader, True)
            te = self._run_epoch(epoch, self.test_loader, False) if self.test_loader else float("nan")
            if self.rank == 0:
                print(f"epoch {epoch + 1}/{self.cfg.max_epochs} train loss {tr:.4f} eval loss {te:.4f}", flush=True)
                if (epoch + 1) % self.cfg.save_every == 0:
                    self._save(epoch + 1)
            dist.barrier()
