"""Character-level dataset of the miniGPT example (reference: ``examples/torch_examples/minigpt/char_dataset.py``)."""
from __future__ import annotations

import torch
from torch.utils.data import Dataset


def generated_corpus(n_words: int = 40000, seed: int = 0) -> str:
    words = ("the quick brown fox jumps over the lazy dog while gradients flow across eight gpus in parallel and every "
             "rank sums its share of the loss before the optimizer takes a step").split()
    g = torch.Generator().manual_seed(seed)
    return " ".join(words[i] for i in torch.randint(0, len(words), (n_words,), generator=g).tolist())


class CharDataset(Dataset):
    """Windows of ``block_size + 1`` characters: input = first ``block_size``, target = the same shifted by one."""

    def __init__(self, path: str | None, block_size: int, truncate: float = 1.0):
        text = open(path, encoding="utf-8").read() if path else generated_corpus()
        text = text[: max(block_size + 2, int(len(text) * truncate))]
        self.chars = sorted(set(text))
        self.stoi = {c: i for i, c in enumerate(self.chars)}
        self.itos = dict(enumerate(self.chars))
        self.block_size, self.vocab_size = block_size, len(self.chars)
        self.data = torch.tensor([self.stoi[c] for c in text], dtype=torch.long)

    def __len__(self) -> int:
        return len(self.data) - self.block_size

    def __getitem__(self, i: int):
        chunk = self.data[i:i + self.block_size + 1]
        return chunk[:-1], chunk[1:]

    def decode(self, ids) -> str:
        return "".join(self.itos[int(i)] for i in ids)
