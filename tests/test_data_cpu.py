"""Data pipeline on CPU: tokenise-strategy registry, chunking, `DatasetProcessor` argument checks, the micro-batch
loader (shapes, shift, positions, epochs, resume, data-rank sharding) and the padded pre-training dataset.
Reference suite: ``tests/data/test_data.py`` (same behaviours, against a fake tokenizer -- the box has no network).
"""
import numpy as np
import pytest
import torch

from scaletorch_b200.data import MicroBatchDataLoader
from scaletorch_b200.data import dataset as DS
from scaletorch_b200.data.dataloader import SyntheticTokens
from scaletorch_b200.data.pretrain_dataset import PretrainDataset


class FakeTokenizer:
    """One token per character (its code point modulo 251); pad id 0."""

    pad_token_id = 0

    def __call__(self, texts, add_special_tokens=False, return_attention_mask=False, return_token_type_ids=False,
                 truncation=False, max_length=None, padding=None, return_tensors=None):
        single = isinstance(texts, str)
        rows = [[1 + ord(c) % 250 for c in t] for t in ([texts] if single else texts)]
        if truncation and max_length:
            rows = [r[:max_length] for r in rows]
        if padding == "max_length" and max_length:
            rows = [r + [0] * (max_length - len(r)) for r in rows]
        if return_tensors == "pt":
            return {"input_ids": torch.tensor(rows)}
        return {"input_ids": rows}


class ListDataset(torch.utils.data.Dataset):
    def __init__(self, rows):
        self.rows = rows

    def __len__(self):
        return len(self.rows)

    def __getitem__(self, i):
        return {"input_ids": self.rows[i]}


# ------------------------------------------------------------------------------------------- strategy registry
def test_builtin_strategies_registered():
    assert DS.get_tokenize_strategy("concat_chunk") is DS.tokenizer_group_text
    assert DS.get_tokenize_strategy("pad_truncate") is DS.tokenizer_pad_truncate


def test_unknown_strategy_raises_with_the_registered_names():
    with pytest.raises(KeyError, match="concat_chunk"):
        DS.get_tokenize_strategy("nope")


def test_register_custom_strategy_both_forms_and_overwrite():
    def f(examples, tokenizer, sequence_length, text_column="text"):
        return {"input_ids": [[7] * (sequence_length + 1)]}

    assert DS.register_tokenize_strategy("unit_custom", f) is f
    assert DS.get_tokenize_strategy("unit_custom") is f

    @DS.register_tokenize_strategy("unit_custom")                 # decorator form overwrites
    def g(examples, tokenizer, sequence_length, text_column="text"):
        return {"input_ids": []}

    assert DS.get_tokenize_strategy("unit_custom") is g
    DS._TOKENIZE_STRATEGIES.pop("unit_custom")


# ------------------------------------------------------------------------------------------- chunking
def test_concat_chunk_basic_and_remainder_dropped():
    out = DS.tokenizer_group_text({"text": ["abcdefghij", "klmnopqrstuvwxyz"]}, FakeTokenizer(), 4)       # 26 tokens, chunk 5
    rows = out["input_ids"]
    assert len(rows) == 5 and all(len(r) == 5 for r in rows)
    flat = [t for r in rows for t in r]
    want = [1 + ord(c) % 250 for c in "abcdefghijklmnopqrstuvwxyz"][:25]
    assert flat == want                                          # documents are concatenated in order, tail dropped


def test_concat_chunk_short_text_returns_empty_and_exact_length_one_chunk():
    assert DS.tokenizer_group_text({"text": ["abc"]}, FakeTokenizer(), 8) == {"input_ids": []}
    out = DS.tokenizer_group_text({"text": ["abcde"]}, FakeTokenizer(), 4)
    assert len(out["input_ids"]) == 1 and len(out["input_ids"][0]) == 5


@pytest.mark.parametrize("bad", [0, -3, 2.5, None])
def test_concat_chunk_rejects_bad_sequence_length(bad):
    with pytest.raises(ValueError, match="sequence_length"):
        DS.tokenizer_group_text({"text": ["abcdef"]}, FakeTokenizer(), bad)


def test_concat_chunk_rejects_missing_column_and_empty_examples():
    with pytest.raises(ValueError, match="column"):
        DS.tokenizer_group_text({"body": ["abcdef"]}, FakeTokenizer(), 2)
    with pytest.raises(ValueError):
        DS.tokenizer_group_text({}, FakeTokenizer(), 2)


def test_pad_truncate_one_row_per_document():
    out = DS.tokenizer_pad_truncate({"text": ["ab", "abcdefghij"]}, FakeTokenizer(), 4)
    assert [len(r) for r in out["input_ids"]] == [5, 5]
    assert out["input_ids"][0][2:] == [0, 0, 0]                  # padded
    assert out["input_ids"][1] == [1 + ord(c) % 250 for c in "abcde"]      # truncated


# ------------------------------------------------------------------------------------------- DatasetProcessor
@pytest.mark.parametrize("bad", ["", "   ", 123, None])
def test_processor_rejects_bad_tokenizer_name(bad):
    with pytest.raises(ValueError):
        DS.DatasetProcessor(tokenizer_name_or_path=bad)


def test_processor_accepts_a_tokenizer_object_and_tokenizes_an_hf_dataset():
    datasets = pytest.importorskip("datasets")
    ds = datasets.Dataset.from_dict({"text": ["abcdefghij", "klmnopqrst"], "meta": [1, 2]})
    proc = DS.DatasetProcessor(tokenizer=FakeTokenizer())
    out = proc.tokenize_dataset(ds, "text", sequence_length=4)
    assert out.column_names == ["input_ids"] and len(out) == 4 and all(len(r) == 5 for r in out["input_ids"])
    with pytest.raises(ValueError, match="column"):
        proc.tokenize_dataset(ds, "body", sequence_length=4)
    with pytest.raises(ValueError, match="sequence_length"):
        proc.tokenize_dataset(ds, "text", sequence_length=0)
    with pytest.raises(ValueError, match="num_proc"):
        proc.tokenize_dataset(ds, "text", sequence_length=4, num_proc=0)
    with pytest.raises(KeyError):
        proc.tokenize_dataset(ds, "text", sequence_length=4, strategy="nope")


def test_load_dataset_from_disk_with_num_samples(tmp_path):
    datasets = pytest.importorskip("datasets")
    datasets.Dataset.from_dict({"text": [f"doc {i}" for i in range(10)]}).save_to_disk(str(tmp_path / "d"))
    ds = DS.DatasetProcessor.load_dataset(str(tmp_path / "d"), num_samples=3)
    assert len(ds) == 3 and ds[0]["text"] == "doc 0"
    dd = datasets.DatasetDict({"train": datasets.Dataset.from_dict({"text": ["a", "b"]}),
                               "validation": datasets.Dataset.from_dict({"text": ["c"]})})
    dd.save_to_disk(str(tmp_path / "dd"))
    assert len(DS.DatasetProcessor.load_dataset(str(tmp_path / "dd"), split="validation")) == 1


def test_export_flat_tokens(tmp_path):
    path = str(tmp_path / "sub" / "tokens.npy")
    n = DS.export_flat_tokens(["ab", "cde"], FakeTokenizer(), path, eos_token_id=9, batch_size=1)
    flat = np.load(path)
    assert n == 7 and flat.dtype == np.int32
    assert flat.tolist() == [1 + ord("a") % 250, 1 + ord("b") % 250, 9, 1 + ord("c") % 250, 1 + ord("d") % 250,
                             1 + ord("e") % 250, 9]
    assert DS.export_flat_tokens([], FakeTokenizer(), str(tmp_path / "e.npy")) == 0


# ------------------------------------------------------------------------------------------- MicroBatchDataLoader
@pytest.mark.parametrize("kw", [dict(micro_batch_size=0), dict(micro_batch_size=-1), dict(micro_batch_size=1.5),
                                dict(sequence_length=0), dict(sequence_length=-8), dict(gradient_accumulation_steps=0),
                                dict(num_samples=0), dict(num_samples=-5), dict(num_workers=-1)])
def test_loader_rejects_bad_arguments(kw):
    base = dict(micro_batch_size=2, sequence_length=8, synthetic=True, vocab_size=50, num_samples=16)
    base.update(kw)
    with pytest.raises(ValueError):
        MicroBatchDataLoader(**base)


def test_loader_needs_vocab_for_synthetic_and_enough_samples():
    with pytest.raises(ValueError, match="vocab_size"):
        MicroBatchDataLoader(2, 8, synthetic=True)
    with pytest.raises(ValueError, match="samples"):
        MicroBatchDataLoader(4, 8, dataset=ListDataset([list(range(9))] * 2))


def test_loader_single_process_attributes():
    dl = MicroBatchDataLoader(2, 8, gradient_accumulation_steps=3, synthetic=True, vocab_size=50, num_samples=16)
    assert dl.micro_batch_size == 2 and dl.sequence_length == 8 and dl.gradient_accumulation_steps == 3
    assert dl.cp_world_size == 1 and dl.cp_rank == 0 and dl.data_world_size == 1 and dl.data_rank == 0
    assert dl.sequence_length_per_rank == 8 and dl.global_batch_size == 6
    assert len(dl) == 8                                          # 16 samples / micro-batch 2, drop_last


def test_collate_shapes_shift_positions_and_hidden_states():
    rows = [list(range(10 * i, 10 * i + 9)) for i in range(4)]   # S + 1 = 9 tokens per sample
    dl = MicroBatchDataLoader(2, 8, dataset=ListDataset(rows))
    b = next(dl)
    assert set(b) == {"input_ids", "target_ids", "position_ids", "hidden_states"} and b["hidden_states"] is None
    assert b["input_ids"].shape == b["target_ids"].shape == b["position_ids"].shape == (2, 8)
    assert b["input_ids"].dtype == torch.long
    assert torch.equal(b["target_ids"][:, :-1], b["input_ids"][:, 1:])            # targets = inputs shifted by one
    assert b["input_ids"][0].tolist() == rows[0][:8] and b["target_ids"][0].tolist() == rows[0][1:]
    assert torch.equal(b["position_ids"], torch.arange(8).expand(2, 8))
    with pytest.raises(ValueError, match="empty"):
        dl.collate_batch([])
    with pytest.raises(ValueError, match="sequence_length \\+ 1"):
        dl.collate_batch([{"input_ids": list(range(8))}])       # one token short: no target for the last position


def test_loader_wraps_around_epochs_and_counts_batches():
    rows = [[i] * 5 for i in range(6)]
    dl = MicroBatchDataLoader(2, 4, dataset=ListDataset(rows))
    firsts = [int(next(dl)["input_ids"][0, 0]) for _ in range(7)]
    assert firsts == [0, 2, 4, 0, 2, 4, 0]                       # 3 batches per epoch, then restart
    assert dl.batches_served == 7 and dl.state_dict()["batches_served"] == 7


def test_loader_resume_reproduces_the_stream_with_shuffle():
    ds = SyntheticTokens(97, 8, num_samples=24, seed=3)
    a = MicroBatchDataLoader(2, 8, dataset=ds, shuffle=True, seed=11)
    stream = [next(a)["input_ids"].clone() for _ in range(20)]   # 12 batches per epoch: crosses an epoch boundary
    b = MicroBatchDataLoader(2, 8, dataset=ds, shuffle=True, seed=11)
    for _ in range(14):
        next(b)
    state = b.state_dict()
    c = MicroBatchDataLoader(2, 8, dataset=ds, shuffle=True, seed=11)
    c.load_state_dict(state)
    for i in range(14, 20):
        assert torch.equal(next(c)["input_ids"], stream[i]), i
    # the two epochs are shuffled differently
    e0 = torch.cat(stream[:12]).tolist()
    e1 = torch.cat(stream[12:20] + [next(a)["input_ids"] for _ in range(4)]).tolist()
    assert sorted(map(tuple, e0)) == sorted(map(tuple, e1)) and e0 != e1


def test_synthetic_tokens_are_deterministic_per_index():
    ds = SyntheticTokens(1000, 16, num_samples=8, seed=5)
    assert len(ds) == 8 and ds[3]["input_ids"].shape == (17,)
    assert torch.equal(ds[3]["input_ids"], ds[3]["input_ids"])
    assert not torch.equal(ds[3]["input_ids"], ds[4]["input_ids"])
    assert not torch.equal(ds[3]["input_ids"], SyntheticTokens(1000, 16, 8, seed=6)[3]["input_ids"])
    assert int(ds[0]["input_ids"].max()) < 1000


# ------------------------------------------------------------------------------------------- PretrainDataset
def test_pretrain_dataset_pads_and_masks():
    ds = PretrainDataset(["abc", "abcdefghijkl"], FakeTokenizer(), max_length=8)
    assert len(ds) == 2
    s = ds[0]
    assert s["input_ids"].shape == s["target_ids"].shape == s["loss_mask"].shape == (7,)
    assert s["loss_mask"].tolist() == [1, 1, 0, 0, 0, 0, 0]       # 3 real tokens -> 2 real targets
    assert s["target_ids"].tolist()[2:] == [-100] * 5
    full = ds[1]
    assert full["loss_mask"].sum().item() == 7 and (full["target_ids"] != -100).all()


def test_loader_position_carries_over_to_another_micro_batch_size():
    ds = SyntheticTokens(97, 8, num_samples=64, seed=3)
    a = MicroBatchDataLoader(4, 8, dataset=ds)
    for _ in range(5):
        next(a)
    state = a.state_dict()
    assert state["samples_consumed"] == 20 and state["micro_batch_size"] == 4 and state["data_world_size"] == 1
    b = MicroBatchDataLoader(2, 8, dataset=ds)                  # resumed with half the micro-batch: 20 samples = 10 batches in
    b.load_state_dict(state)
    assert b.batches_served == 10
    nxt = next(b)["input_ids"]
    assert torch.equal(nxt[0], ds[20]["input_ids"][:8]) and torch.equal(nxt[1], ds[21]["input_ids"][:8])
    c = MicroBatchDataLoader(4, 8, dataset=ds)                  # same layout: the batch counter is used as before
    c.load_state_dict({"epoch": 0, "batches_served": 5})
    assert torch.equal(next(c)["input_ids"][0], ds[20]["input_ids"][:8])
