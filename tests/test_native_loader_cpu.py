"""The C++ micro-batch producer (csrc/runtime/token_loader.cpp) against its pure-Python oracle."""
import pytest
import torch

from scaletorch_b200.data.native_loader import (NativeTokenLoader, epoch_permutation, native_loader_available,
                                                reference_batches)

needs_lib = pytest.mark.skipif(not native_loader_available(), reason="native library not built (python -m scaletorch_b200._build)")


def test_epoch_permutation_is_a_permutation_and_deterministic():
    p0, p1 = epoch_permutation(101, 7, 0, True), epoch_permutation(101, 7, 1, True)
    assert sorted(p0) == list(range(101)) and sorted(p1) == list(range(101)) and p0 != p1
    assert p0 == epoch_permutation(101, 7, 0, True)
    assert epoch_permutation(5, 7, 3, False) == [0, 1, 2, 3, 4]


@needs_lib
@pytest.mark.parametrize("cp_size,zigzag,data_world,shuffle", [(1, False, 1, True), (2, False, 2, True), (2, True, 1, True),
                                                               (4, True, 2, False)])
def test_native_loader_matches_oracle(cp_size, zigzag, data_world, shuffle):
    torch.manual_seed(0)
    S, mbs = 32, 3
    tokens = torch.randint(0, 50000, ((S + 1) * 41 + 5,), dtype=torch.int32)
    for cp_rank in range(cp_size):
        for data_rank in range(data_world):
            kw = dict(cp_rank=cp_rank, cp_size=cp_size, zigzag=zigzag, data_rank=data_rank, data_world=data_world,
                      shuffle=shuffle, seed=1234)
            loader = NativeTokenLoader(tokens, S, mbs, n_slots=3, n_threads=3, **kw)
            n = 2 * loader.batches_per_epoch + 2           # crosses two epoch boundaries (new permutation each epoch)
            ref = reference_batches(tokens, S, mbs, n, **kw)
            for b in range(n):
                got = next(loader)
                assert torch.equal(got["input_ids"], ref[b][0]), (cp_rank, data_rank, b)
                assert torch.equal(got["target_ids"], ref[b][1])
                assert got["position_ids"].shape == (mbs, S // cp_size)
            loader.close()


@needs_lib
def test_native_loader_feeds_the_train_step():
    from types import SimpleNamespace

    from scaletorch_b200.models import Qwen3
    from scaletorch_b200.parallel.context import ParallelContext
    from scaletorch_b200.trainer import train_step

    cfg = SimpleNamespace(vocab_size=512, hidden_size=64, num_attention_heads=2, num_key_value_heads=1, head_dim=32,
                          max_position_embeddings=128, num_hidden_layers=1, intermediate_size=128, rms_norm_eps=1e-6,
                          rope_theta=1e4, tie_word_embeddings=True, model_type="qwen3")
    model = Qwen3(cfg, ParallelContext())
    tokens = torch.randint(0, 512, (65 * 20,), dtype=torch.int32)
    loader = NativeTokenLoader(tokens, 64, 2, gradient_accumulation_steps=2)
    loss = train_step(model, loader, torch.device("cpu"), torch.float32, 2)
    assert torch.isfinite(loss) and loader.batches_served == 2
    loader.close()


@needs_lib
def test_native_loader_rejects_bad_layouts():
    tokens = torch.zeros(1000, dtype=torch.int32)
    with pytest.raises(ValueError):
        NativeTokenLoader(tokens, 30, 2, cp_size=4, zigzag=True)      # 30 % 8 != 0
    with pytest.raises(ValueError):
        NativeTokenLoader(tokens, 64, 64)                               # fewer samples than one batch


def test_export_flat_tokens_roundtrip(tmp_path):
    import numpy as np

    from scaletorch_b200.data.dataset import export_flat_tokens

    class FakeTok:
        def __call__(self, texts, **kw):
            return {"input_ids": [[ord(c) for c in t] for t in texts]}

    path = str(tmp_path / "sub" / "tokens.npy")
    n = export_flat_tokens(["ab", "c", "def"], FakeTok(), path, eos_token_id=0, batch_size=2)
    arr = np.load(path)
    assert n == 9 and arr.dtype == np.int32
    assert arr.tolist() == [97, 98, 0, 99, 0, 100, 101, 102, 0]
    assert export_flat_tokens([], FakeTok(), path) == 0
