"""Held-out loss / perplexity of a model under any data-parallel width.

    python tools/evaluate.py --model_name_or_path export/step300 --load_pretrained_weights true --token_file val.npy \
        --sequence_length 2048 --micro_batch_size 4 [--max_batches 100]
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 tools/evaluate.py … --data_parallel_size 8

Takes the trainer's flags (``ScaleTorchArguments``) plus ``--max_batches``.  Data: ``--token_file`` (flat int32 token ids,
``.npy`` or raw ``.bin``; windows of ``S + 1`` tokens, no overlap), a dataset through ``--dataset_name`` /
``--tokenizer_name_or_path``, or ``--synthetic_data true``.  Every data-parallel rank evaluates its shard of the windows;
token counts and summed losses meet in ONE ``all_reduce_dict`` at the end (``scaletorch_b200.dist``), rank 0 prints one
JSON line: ``{"loss", "perplexity", "tokens", "batches", "tokens_per_s"}``.  Tensor / context parallel layouts work as in
training (the loss is already reduced over them); pipeline-parallel evaluation is not offered.
"""
from __future__ import annotations

import json
import math
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from transformers import HfArgumentParser  # noqa: E402

from scaletorch_b200 import dist as D  # noqa: E402
from scaletorch_b200.parallel.process_group import process_group_manager as pgm  # noqa: E402
from scaletorch_b200.trainer import (ScaleTorchArguments, cleanup_distributed, create_model, get_dtype,  # noqa: E402
                                     initialize_distributed_training, validate_config)


class TokenWindows(torch.utils.data.Dataset):
    """Non-overlapping windows of ``S + 1`` ids from a flat token array."""

    def __init__(self, tokens: np.ndarray, sequence_length: int):
        self.tokens, self.chunk = tokens, sequence_length + 1
        self.n = len(tokens) // self.chunk
        if self.n == 0:
            raise ValueError(f"token file holds {len(tokens)} tokens; one window needs {self.chunk}")

    def __len__(self) -> int:
        return self.n

    def __getitem__(self, i: int):
        w = self.tokens[i * self.chunk:(i + 1) * self.chunk]
        return {"input_ids": torch.from_numpy(np.asarray(w, dtype=np.int64))}


def load_tokens(path: str) -> np.ndarray:
    return np.load(path, mmap_mode="r") if path.endswith(".npy") else np.memmap(path, dtype=np.int32, mode="r")


@torch.no_grad()
def evaluate(model, loader, device, max_batches: int | None = None) -> dict:
    """Sum of token losses and token count over this rank's batches (one pass over the loader's epoch)."""
    raw = model.module if hasattr(model, "module") else model
    raw.eval()
    n_batches = len(loader) if max_batches is None else min(len(loader), max_batches)
    loss_sum = torch.zeros((), dtype=torch.float64, device=device)
    tokens = 0
    t0 = time.time()
    for _ in range(n_batches):
        b = next(loader)
        x, y = b["input_ids"].to(device, non_blocking=True), b["target_ids"].to(device, non_blocking=True)
        n = int((y != -100).sum())
        loss_sum += raw.loss(x, y).double() * n              # mean over the batch's tokens -> sum
        tokens += n
    if device.type == "cuda":
        torch.cuda.synchronize()
    return {"loss_sum": loss_sum, "tokens": tokens, "batches": n_batches, "seconds": time.time() - t0}


def main(argv=None) -> dict | None:
    argv = list(sys.argv[1:] if argv is None else argv)
    max_batches = None
    if "--max_batches" in argv:
        i = argv.index("--max_batches")
        max_batches = int(argv[i + 1])
        del argv[i:i + 2]
    (config,) = HfArgumentParser(ScaleTorchArguments).parse_args_into_dataclasses(args=argv)
    if config.pipeline_parallel_size > 1:
        raise SystemExit("tools/evaluate.py: evaluate with pipeline_parallel_size = 1 (convert the checkpoint first: tools/convert_checkpoint.py)")
    dtype = get_dtype(config)
    _, rank, world, device = initialize_distributed_training(config)
    validate_config(config, world)
    try:
        from scaletorch_b200.utils.misc import set_all_seed

        set_all_seed(config.seed)                            # random-init models (no checkpoint) are the same under every layout
        model, mc = create_model(config, dtype, device)
        from scaletorch_b200.data import MicroBatchDataLoader

        dataset = TokenWindows(load_tokens(config.token_file), config.sequence_length) if config.token_file else None
        loader = MicroBatchDataLoader(
            config.micro_batch_size, config.sequence_length, dataset_name=config.dataset_name,
            tokenizer_name_or_path=config.tokenizer_name_or_path, num_samples=config.num_samples, split=config.split,
            subset_name=config.subset_name, pin_memory=config.pin_memory, synthetic=config.synthetic_data and dataset is None,
            vocab_size=mc.vocab_size, seed=config.seed, dataset=dataset, num_workers=0)
        part = evaluate(model, loader, device, max_batches)
        stats = {"loss_sum": part["loss_sum"].reshape(1), "tokens": torch.tensor([float(part["tokens"])], dtype=torch.float64, device=device),
                 "batches": torch.tensor([float(part["batches"])], dtype=torch.float64, device=device)}
        if pgm and pgm.grad_sync_world_size > 1:
            # data ranks hold disjoint windows, context-parallel ranks disjoint chunks of the same windows: every token is
            # counted exactly once in the sum over the CP x DP (x EP) group; tensor-parallel ranks hold identical values
            batches = stats.pop("batches")
            D.all_reduce_dict(stats, op="sum", group=pgm.grad_sync_group)
            stats["batches"] = batches
        tokens = float(stats["tokens"].item())
        loss = float(stats["loss_sum"].item()) / max(tokens, 1.0)
        out = {"loss": loss, "perplexity": math.exp(min(loss, 50.0)), "tokens": int(tokens), "batches": int(stats["batches"].item()),
               "tokens_per_s": tokens / max(part["seconds"], 1e-9), "model": config.model_name_or_path, "world_size": world}
        if rank == 0:
            print(json.dumps(out), flush=True)
        return out
    finally:
        cleanup_distributed()


if __name__ == "__main__":
    main()
