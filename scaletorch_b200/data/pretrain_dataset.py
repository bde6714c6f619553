"""Padded per-sample pre-training dataset (reference: ``scaletorch/data/pretrain_dataset.py:1-107``)."""
from __future__ import annotations

import torch
from torch.utils.data import Dataset


class PretrainDataset(Dataset):
    """Each text becomes one sample of ``max_length`` ids; the loss mask excludes padding."""

    def __init__(self, texts: list[str], tokenizer, max_length: int = 512):
        self.texts, self.tokenizer, self.max_length = texts, tokenizer, max_length

    def __len__(self) -> int:
        return len(self.texts)

    def __getitem__(self, idx: int):
        enc = self.tokenizer(str(self.texts[idx]), max_length=self.max_length, padding="max_length", truncation=True,
                             return_tensors="pt")
        ids = enc["input_ids"].squeeze(0)
        pad = self.tokenizer.pad_token_id
        mask = (ids != pad) if pad is not None else torch.ones_like(ids, dtype=torch.bool)
        x, y = ids[:-1].clone(), ids[1:].clone()
        y[~mask[1:]] = -100
        return {"input_ids": x, "target_ids": y, "loss_mask": mask[1:].long()}
