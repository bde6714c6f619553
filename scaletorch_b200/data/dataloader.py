"""``MicroBatchDataLoader``: micro-batches with CP slicing, DP sharding and host-side prefetch.

Reference: ``scaletorch/data/dataloader.py:35-292``.  Same constructor vocabulary and the same batch dict
(``input_ids``, ``target_ids``, ``position_ids``, ``hidden_states``); iteration restarts with
``set_epoch`` at exhaustion.  Differences: the sampler is keyed on ``pgm.data_rank`` so EP ranks read
distinct samples unless ``ep_replicate_data``; batches are assembled in *pinned* host
memory and copied with ``non_blocking=True``; a synthetic token source (random ids of the configured
shape, regenerated every step on the host) serves the benchmarks on boxes without datasets.
"""
from __future__ import annotations

from typing import Any

import torch
from torch.utils.data import DataLoader, DistributedSampler

from ..parallel.process_group import process_group_manager as pgm


class SyntheticTokens(torch.utils.data.Dataset):
    """``num_samples`` random sequences of ``sequence_length + 1`` ids (deterministic per index)."""

    def __init__(self, vocab_size: int, sequence_length: int, num_samples: int = 1 << 20, seed: int = 0):
        self.vocab_size, self.sequence_length, self.num_samples, self.seed = vocab_size, sequence_length, num_samples, seed

    def __len__(self) -> int:
        return self.num_samples

    def __getitem__(self, idx: int):
        g = torch.Generator().manual_seed(self.seed * 1_000_003 + idx)
        return {"input_ids": torch.randint(0, self.vocab_size, (self.sequence_length + 1,), generator=g)}


class TokenWindows(torch.utils.data.Dataset):
    """Non-overlapping windows of ``sequence_length + 1`` ids from a flat token array (held-out evaluation data)."""

    def __init__(self, tokens, sequence_length: int):
        self.tokens, self.chunk = tokens, sequence_length + 1
        self.n = len(tokens) // self.chunk
        if self.n == 0:
            raise ValueError(f"token file holds {len(tokens)} tokens; one window needs {self.chunk}")

    def __len__(self) -> int:
        return self.n

    def __getitem__(self, i: int):
        import numpy as np

        w = self.tokens[i * self.chunk:(i + 1) * self.chunk]
        return {"input_ids": torch.from_numpy(np.asarray(w, dtype=np.int64))}


def load_tokens(path: str):
    """Memory-map a flat token file: ``.npy`` (any integer dtype) or raw int32 ``.bin``."""
    import numpy as np

    return np.load(path, mmap_mode="r") if path.endswith(".npy") else np.memmap(path, dtype=np.int32, mode="r")


class MicroBatchDataLoader(DataLoader):
    """Loader of ``{input_ids, target_ids, position_ids}`` micro-batches: a ``DistributedSampler`` over the DATA ranks, targets shifted by one token, each sample cut to this context-parallel rank's slice (contiguous or zig-zag), one batch of software prefetch, automatic epoch roll-over, and a resumable position (``state_dict`` / ``load_state_dict``)."""
    def __init__(self, micro_batch_size: int, sequence_length: int, dataset_name: str | None = None,
                 tokenizer_name_or_path: str | None = None, gradient_accumulation_steps: int = 1, device=None,
                 num_workers: int = 0, num_proc: int = 1, num_samples: int | None = None, split: str = "train",
                 subset_name: str | None = None, pin_memory: bool = True, shuffle: bool = False, seed: int = 0,
                 synthetic: bool = False, vocab_size: int | None = None, dataset=None, tokenizer=None,
                 text_column_name: str = "text", strategy: str = "concat_chunk"):
        # argument checks of the reference loader (``dataloader.py:60-105``): fail at construction, not in a worker
        for name, v in (("micro_batch_size", micro_batch_size), ("sequence_length", sequence_length),
                        ("gradient_accumulation_steps", gradient_accumulation_steps)):
            if not isinstance(v, int) or isinstance(v, bool) or v < 1:
                raise ValueError(f"{name} must be a positive integer, got {v!r}")
        if num_samples is not None and (not isinstance(num_samples, int) or num_samples < 1):
            raise ValueError(f"num_samples must be a positive integer or None, got {num_samples!r}")
        if not isinstance(num_workers, int) or num_workers < 0:
            raise ValueError(f"num_workers must be a non-negative integer, got {num_workers!r}")
        self.micro_batch_size = micro_batch_size
        self.sequence_length = sequence_length
        self.gradient_accumulation_steps = gradient_accumulation_steps
        self.cp_world_size = pgm.cp_world_size if pgm else 1
        self.cp_rank = pgm.cp_rank if pgm else 0
        self.data_world_size = pgm.data_world_size if pgm else 1
        self.data_rank = pgm.data_rank if pgm else 0
        if sequence_length % self.cp_world_size != 0:
            raise ValueError(f"sequence_length ({sequence_length}) must be divisible by cp size ({self.cp_world_size})")
        self.sequence_length_per_rank = sequence_length // self.cp_world_size
        self.global_batch_size = micro_batch_size * gradient_accumulation_steps * self.data_world_size
        if dataset is None:
            if synthetic:
                if vocab_size is None:
                    raise ValueError("synthetic data needs vocab_size")
                dataset = SyntheticTokens(vocab_size, sequence_length, num_samples or (1 << 20), seed)
            else:
                from .dataset import DatasetProcessor

                proc = DatasetProcessor(tokenizer_name_or_path, tokenizer)
                raw = proc.load_dataset(dataset_name, split, subset_name, num_samples)
                dataset = proc.tokenize_dataset(raw, text_column_name, sequence_length, num_proc, strategy)
        self.tokenized_dataset = dataset
        if len(dataset) < self.data_world_size * micro_batch_size:
            raise ValueError(f"dataset has {len(dataset)} samples; need at least {self.data_world_size * micro_batch_size}")
        self.sampler_ = DistributedSampler(dataset, num_replicas=self.data_world_size, rank=self.data_rank,
                                           shuffle=shuffle, seed=seed, drop_last=True)
        use_pin = bool(pin_memory) and torch.cuda.is_available()
        super().__init__(dataset, batch_size=micro_batch_size, collate_fn=self.collate_batch, pin_memory=use_pin,
                         num_workers=num_workers, sampler=self.sampler_, shuffle=False, drop_last=True,
                         persistent_workers=num_workers > 0, prefetch_factor=2 if num_workers > 0 else None)
        self._iterator = None
        self._epoch = 0
        self._prefetched = None
        self.batches_served = 0

    def collate_batch(self, batch: list[dict[str, Any]]) -> dict[str, Any]:
        """Stack → CP-slice ``[cp_rank·S/cp, +S/cp)`` → shifted targets and absolute positions."""
        if not batch:
            raise ValueError("collate_batch got an empty batch")
        ids = torch.stack([torch.as_tensor(b["input_ids"], dtype=torch.long) for b in batch])  # [mbs, S+1]
        if ids.shape[1] < self.sequence_length + 1:
            raise ValueError(f"samples hold {ids.shape[1]} tokens; the loader needs sequence_length + 1 = "
                             f"{self.sequence_length + 1} (targets are the inputs shifted by one)")
        if self.cp_world_size > 1:
            from ..parallel.context_parallel import zigzag

            if zigzag.enabled():   # load-balanced causal CP: chunks cp_rank and 2·cp−1−cp_rank of 2·cp
                S = self.sequence_length_per_rank * self.cp_world_size
                input_ids = zigzag.take(ids, S, self.cp_world_size, self.cp_rank).contiguous()
                target_ids = zigzag.take(ids, S, self.cp_world_size, self.cp_rank, shift=1).contiguous()
                pos = zigzag.local_positions(self.sequence_length_per_rank, self.cp_world_size, self.cp_rank)
                position_ids = pos.unsqueeze(0).expand(ids.shape[0], -1).contiguous()
                return {"input_ids": input_ids, "target_ids": target_ids, "position_ids": position_ids, "hidden_states": None}
        lo = self.cp_rank * self.sequence_length_per_rank
        hi = lo + self.sequence_length_per_rank
        input_ids = ids[:, lo:hi].contiguous()
        target_ids = ids[:, lo + 1:hi + 1].contiguous()
        position_ids = torch.arange(lo, hi, dtype=torch.long).unsqueeze(0).expand(ids.shape[0], -1).contiguous()
        return {"input_ids": input_ids, "target_ids": target_ids, "position_ids": position_ids, "hidden_states": None}

    def __iter__(self):
        if self._iterator is None:
            self.sampler_.set_epoch(self._epoch)
            self._iterator = super().__iter__()
        return self

    def _next_raw(self):
        if self._iterator is None:
            self.__iter__()
        try:
            return next(self._iterator)
        except StopIteration:
            self._epoch += 1
            self.sampler_.set_epoch(self._epoch)
            self._iterator = super().__iter__()
            return next(self._iterator)

    def __next__(self):
        """One-batch software prefetch: the batch returned now was assembled during the previous step."""
        if self._prefetched is None:
            self._prefetched = self._next_raw()
        out = self._prefetched
        self._prefetched = self._next_raw()
        self.batches_served += 1
        return out

    def state_dict(self) -> dict:
        """Position of the sample stream.  ``samples_consumed`` (over all data ranks) is layout-independent: a checkpoint
        converted to another layout (``tools/convert_checkpoint.py``) resumes at the same place in the data."""
        return {"epoch": self._epoch, "batches_served": self.batches_served, "micro_batch_size": self.micro_batch_size,
                "data_world_size": self.data_world_size,
                "samples_consumed": self.batches_served * self.micro_batch_size * self.data_world_size}

    def load_state_dict(self, state: dict) -> None:
        """Resume the sample stream where a checkpoint left it: same epoch (= same shuffle) and position, so a resumed run
        sees exactly the batches the uninterrupted run would have seen (the reference restarts the sampler on resume,
        SURVEY §5.4).  The skipped batches of the current epoch are drawn and discarded."""
        served = int(state.get("batches_served", 0))
        same_layout = (state.get("micro_batch_size", self.micro_batch_size) == self.micro_batch_size
                       and state.get("data_world_size", self.data_world_size) == self.data_world_size)
        if not same_layout and "samples_consumed" in state:
            # saved under another micro-batch size / number of data ranks: continue after the same number of samples
            served = int(state["samples_consumed"]) // (self.micro_batch_size * self.data_world_size)
        per_epoch = max(1, len(self.sampler_) // self.micro_batch_size)
        self._epoch, skip = divmod(served, per_epoch)
        self.sampler_.set_epoch(self._epoch)
        self._iterator = super().__iter__()
        self._prefetched = None
        for _ in range(skip):
            next(self._iterator)
        self.batches_served = served
