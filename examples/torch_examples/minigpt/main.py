"""miniGPT entry point: YAML config + ``section.key=value`` overrides → dataset, model, optimiser, DDP trainer
(reference: ``examples/torch_examples/minigpt/main.py``, which uses hydra; plain YAML here, no extra dependency).

    torchrun --standalone --local-addr 127.0.0.1 --nproc-per-node 2 examples/torch_examples/minigpt/main.py \
        trainer.max_epochs=2 trainer.max_steps_per_epoch=50
"""
from __future__ import annotations

import os
import sys

import torch
import torch.distributed as dist
import yaml
from torch.utils.data import random_split

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, ".."))
from char_dataset import CharDataset  # noqa: E402
from common import setup  # noqa: E402
from model import GPT, GPTConfig  # noqa: E402
from trainer import Trainer, TrainerConfig  # noqa: E402


def load_config(argv: list[str]) -> dict:
    path = os.path.join(HERE, "gpt2_train_cfg.yaml")
    overrides = []
    for a in argv:
        if a.startswith("--config="):
            path = a.split("=", 1)[1]
        elif "=" in a:
            overrides.append(a)
    cfg = yaml.safe_load(open(path))
    for o in overrides:
        key, val = o.split("=", 1)
        section, name = key.split(".", 1)
        cfg[section][name] = yaml.safe_load(val)
    return cfg


def main() -> None:
    cfg = load_config(sys.argv[1:])
    rank, world, device = setup()
    data = CharDataset(cfg["data"]["path"], cfg["data"]["block_size"], cfg["data"]["truncate"])
    n_train = int(len(data) * cfg["data"]["train_split"])
    train_set, test_set = random_split(data, [n_train, len(data) - n_train], generator=torch.Generator().manual_seed(0))
    torch.manual_seed(0)
    model = GPT(GPTConfig(vocab_size=data.vocab_size, block_size=cfg["data"]["block_size"], **cfg["model"]))
    opt = torch.optim.AdamW(model.optimizer_groups(cfg["optimizer"]["weight_decay"]), lr=cfg["optimizer"]["learning_rate"],
                            betas=(0.9, 0.95))
    trainer = Trainer(TrainerConfig(**cfg["trainer"]), model, opt, train_set, test_set, device, rank, world)
    trainer.train()
    if rank == 0:
        raw = trainer.model.module
        raw.eval()
        prompt = torch.tensor([[data.stoi[c] for c in "the "]], device=device)
        print("sample:", repr(data.decode(raw.generate(prompt, 80, temperature=0.8, top_k=10)[0].tolist())))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
