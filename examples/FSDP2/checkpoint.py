"""Distributed checkpointing for sharded (DTensor) models with ``torch.distributed.checkpoint`` (DCP): every rank writes
its own shards, and a checkpoint can be loaded under a different world size / sharding (topics of the reference's
``examples/FSDP2/checkpoint.py:43-70,234-313``)."""
from __future__ import annotations

import os

import torch.distributed as dist
import torch.distributed.checkpoint as dcp
from torch.distributed.checkpoint.state_dict import get_state_dict, set_state_dict
from torch.distributed.checkpoint.stateful import Stateful


class TrainState(Stateful):
    """Wraps model + optimiser so DCP calls the DTensor-aware state-dict getters/setters."""

    def __init__(self, model, optimizer):
        self.model, self.optimizer = model, optimizer

    def state_dict(self):
        m, o = get_state_dict(self.model, self.optimizer)
        return {"model": m, "optim": o}

    def load_state_dict(self, sd):
        set_state_dict(self.model, self.optimizer, model_state_dict=sd["model"], optim_state_dict=sd["optim"])


def save_checkpoint(model, optimizer, directory: str, step: int) -> None:
    path = os.path.join(directory, f"step_{step}")
    dcp.save({"app": TrainState(model, optimizer)}, checkpoint_id=path)
    if dist.get_rank() == 0:
        with open(os.path.join(directory, "latest"), "w") as f:
            f.write(str(step))
        print(f"saved DCP checkpoint {path}")


def load_checkpoint(model, optimizer, directory: str) -> int:
    latest = os.path.join(directory, "latest")
    if not os.path.exists(latest):
        return 0
    step = int(open(latest).read().strip())
    dcp.load({"app": TrainState(model, optimizer)}, checkpoint_id=os.path.join(directory, f"step_{step}"))
    if dist.get_rank() == 0:
        print(f"resumed from step {step}")
    return step
