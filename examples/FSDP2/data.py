"""Token data pipeline of the Llama-2 FSDP2 examples (reference: ``examples/FSDP2/data.py``).

Three sources, all yielding ``(input_ids [B,S], labels [B,S])`` for next-token prediction:
  * ``--data synthetic``            random tokens (no files needed; deterministic per rank and step)
  * ``--data tokens --path x.bin``  a flat int32 token file (``tools/prepare_tokens.py`` writes one from any text corpus)
  * ``--data text --path x.txt``    a UTF-8 text file encoded at the byte level (vocabulary 256 + <bos>)
Every rank reads a disjoint, strided set of windows; ``state_dict`` / ``load_state_dict`` make the position resumable.
"""
from __future__ import annotations

import numpy as np
import torch


class TokenWindows:
    def __init__(self, tokens: np.ndarray | None, vocab: int, batch: int, seq: int, rank: int, world: int, seed: int = 0):
        self.tokens, self.vocab, self.batch, self.seq = tokens, vocab, batch, seq
        self.rank, self.world, self.seed = rank, world, seed
        self.step = 0
        if tokens is not None:
            self.n_windows = (len(tokens) - 1) // seq
            if self.n_windows < batch * world:
                raise ValueError(f"corpus has {self.n_windows} windows of {seq} tokens; need at least {batch * world}")

    def __iter__(self):
        return self

    def __next__(self):
        g = torch.Generator().manual_seed(self.seed * 1_000_003 + self.step)
        if self.tokens is None:
            g.manual_seed(self.seed * 1_000_003 + self.step * self.world + self.rank)
            tok = torch.randint(0, self.vocab, (self.batch, self.seq + 1), generator=g)
        else:
            # one global permutation draw per step (same on every rank), rank r takes its slice of it
            idx = torch.randperm(self.n_windows, generator=g)[: self.batch * self.world]
            mine = idx[self.rank * self.batch:(self.rank + 1) * self.batch].tolist()
            tok = torch.stack([torch.from_numpy(self.tokens[i * self.seq:i * self.seq + self.seq + 1].astype(np.int64)) for i in mine])
        self.step += 1
        return tok[:, :-1].contiguous(), tok[:, 1:].contiguous()

    def state_dict(self):
        return {"step": self.step, "seed": self.seed}

    def load_state_dict(self, sd):
        self.step, self.seed = int(sd["step"]), int(sd["seed"])


def build_loader(kind: str, path: str | None, vocab: int, batch: int, seq: int, rank: int, world: int, seed: int = 0):
    if kind == "synthetic":
        return TokenWindows(None, vocab, batch, seq, rank, world, seed)
    if kind == "tokens":
        tokens = np.memmap(path, dtype=np.int32, mode="r")
        return TokenWindows(tokens, vocab, batch, seq, rank, world, seed)
    if kind == "text":
        raw = np.frombuffer(open(path, "rb").read(), dtype=np.uint8).astype(np.int32)
        return TokenWindows(raw, 257, batch, seq, rank, world, seed)
    raise ValueError(f"unknown data source {kind}")
