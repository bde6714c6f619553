"""Dataset tokenisation / chunking with a strategy registry (reference: ``scaletorch/data/dataset.py:25-489``).

``concat_chunk``: concatenate all token ids and cut chunks of ``sequence_length + 1`` (input/target shift
happens in the loader).  The tokenizer is built on rank 0 and broadcast as an object, the dataset comes
from ``load_from_disk`` when ``dataset_name`` is a directory, else ``load_dataset``.
"""
from __future__ import annotations

import os
from typing import Any, Callable

import numpy as np

from ..dist import broadcast_object_list, get_rank, is_distributed
from ..utils.logger_utils import get_logger

logger = get_logger(__name__)

TokenizeFn = Callable[..., dict[str, list]]
_TOKENIZE_STRATEGIES: dict[str, TokenizeFn] = {}


def register_tokenize_strategy(name: str, fn: TokenizeFn | None = None):
    """Register ``fn`` as the tokenize strategy ``name`` (usable as a decorator when ``fn`` is omitted)."""
    def _reg(f: TokenizeFn) -> TokenizeFn:
        _TOKENIZE_STRATEGIES[name] = f
        return f

    return _reg(fn) if fn is not None else _reg


def get_tokenize_strategy(name: str) -> TokenizeFn:
    """Look up a registered tokenize strategy; ``KeyError`` lists the known names."""
    if name not in _TOKENIZE_STRATEGIES:
        raise KeyError(f"unknown tokenize strategy '{name}'; registered: {sorted(_TOKENIZE_STRATEGIES)}")
    return _TOKENIZE_STRATEGIES[name]


@register_tokenize_strategy("concat_chunk")
def tokenizer_group_text(examples: dict[str, list], tokenizer: Any, sequence_length: int, text_column: str = "text"):
    """Concatenates the batch's tokens and emits chunks of ``sequence_length + 1`` ids."""
    if not isinstance(sequence_length, int) or sequence_length < 1:
        raise ValueError("sequence_length must be a positive integer")
    if not examples or text_column not in examples:
        raise ValueError(f"examples must hold a '{text_column}' column")
    enc = tokenizer(examples[text_column], add_special_tokens=False, return_attention_mask=False,
                    return_token_type_ids=False)
    ids = np.fromiter((t for seq in enc["input_ids"] for t in seq), dtype=np.int64)
    chunk = sequence_length + 1
    n = (len(ids) // chunk) * chunk
    if n == 0:
        return {"input_ids": []}
    return {"input_ids": ids[:n].reshape(-1, chunk).tolist()}


@register_tokenize_strategy("pad_truncate")
def tokenizer_pad_truncate(examples: dict[str, list], tokenizer: Any, sequence_length: int, text_column: str = "text"):
    """One sample per document, padded / truncated to ``sequence_length + 1``."""
    enc = tokenizer(examples[text_column], add_special_tokens=False, truncation=True, max_length=sequence_length + 1,
                    padding="max_length", return_attention_mask=False)
    return {"input_ids": enc["input_ids"]}


class DatasetProcessor:
    """Tokenizer construction (built on rank 0 and broadcast), dataset loading (``load_from_disk`` for a directory, ``load_dataset`` otherwise) and tokenisation with a registered strategy through ``datasets.map``."""
    def __init__(self, tokenizer_name_or_path: str | None = None, tokenizer: Any = None, device=None):
        if tokenizer is None and tokenizer_name_or_path is None:
            raise ValueError("either tokenizer or tokenizer_name_or_path must be given")
        if tokenizer is None and (not isinstance(tokenizer_name_or_path, str) or not tokenizer_name_or_path.strip()):
            raise ValueError(f"tokenizer_name_or_path must be a non-empty string, got {tokenizer_name_or_path!r}")
        self.tokenizer = tokenizer if tokenizer is not None else self.create_tokenizer(tokenizer_name_or_path)

    @staticmethod
    def create_tokenizer(name_or_path: str):
        """Builds the tokenizer on rank 0 and broadcasts the object (reference ``dataset.py:153-198``)."""
        from transformers import AutoTokenizer

        if not is_distributed():
            return AutoTokenizer.from_pretrained(name_or_path)
        holder = [AutoTokenizer.from_pretrained(name_or_path) if get_rank() == 0 else None]
        broadcast_object_list(holder, src=0)
        return holder[0]

    @staticmethod
    def load_dataset(dataset_name: str, split: str = "train", subset_name: str | None = None, num_samples: int | None = None):
        import datasets

        if os.path.isdir(dataset_name):
            ds = datasets.load_from_disk(dataset_name)
            if isinstance(ds, datasets.DatasetDict):
                ds = ds[split]
        else:
            ds = datasets.load_dataset(dataset_name, subset_name, split=split)
        if num_samples is not None:
            ds = ds.select(range(min(num_samples, len(ds))))
        return ds

    def tokenize_dataset(self, dataset, text_column_name: str = "text", sequence_length: int = 1024, num_proc: int = 1,
                         strategy: str = "concat_chunk"):
        if not isinstance(sequence_length, int) or sequence_length < 1:
            raise ValueError("sequence_length must be a positive integer")
        if not isinstance(num_proc, int) or num_proc < 1:
            raise ValueError("num_proc must be a positive integer")
        if text_column_name not in dataset.column_names:
            raise ValueError(f"column '{text_column_name}' not found in dataset columns {dataset.column_names}")
        fn = get_tokenize_strategy(strategy)
        return dataset.map(
            lambda ex: fn(ex, self.tokenizer, sequence_length, text_column_name), batched=True,
            remove_columns=dataset.column_names, num_proc=num_proc if num_proc > 1 else None,
            desc=f"tokenize[{strategy}] S={sequence_length}")


def export_flat_tokens(texts, tokenizer: Any, path: str, eos_token_id: int | None = None, batch_size: int = 1000) -> int:
    """Tokenises ``texts`` (any iterable of strings) into ONE flat int32 array -- documents separated by ``eos_token_id``
    when given -- and saves it as ``.npy``: the input format of the native loader (``data/native_loader.py``,
    ``tools/train.py --native_loader true --token_file …``).  Same token stream as the ``concat_chunk`` strategy, without
    materialising per-sample rows.  Returns the number of tokens written."""
    parts: list[np.ndarray] = []
    batch: list[str] = []

    def flush():
        if not batch:
            return
        enc = tokenizer(batch, add_special_tokens=False, return_attention_mask=False, return_token_type_ids=False)
        for seq in enc["input_ids"]:
            parts.append(np.asarray(seq, dtype=np.int32))
            if eos_token_id is not None:
                parts.append(np.asarray([eos_token_id], dtype=np.int32))
        batch.clear()

    for t in texts:
        batch.append(str(t))
        if len(batch) >= batch_size:
            flush()
    flush()
    flat = np.concatenate(parts) if parts else np.zeros(0, dtype=np.int32)
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    np.save(path, flat)
    return int(flat.size)
