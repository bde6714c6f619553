"""``utils/sharding.py::merge_shards`` is the inverse of ``shard_tensor`` for every layout, and
``tools/convert_checkpoint.py`` turns the per-(tp, pp, ep) training checkpoints of a step into an HF export or an
unsharded training checkpoint that loads back into the same function (the reference has no way out of a layout,
``scaletorch/utils/checkpoint.py:467-560``)."""
import importlib.util
import os
import re
import shutil

import pytest
import torch

from dist_utils import TinyCfg, rel_err
from scaletorch_b200.models import Llama, Qwen3, Qwen3MoE
from scaletorch_b200.parallel.context import ParallelContext
from scaletorch_b200.parallel.pipeline_parallel.pipeline_parallel import distribute_layers
from scaletorch_b200.utils.sharding import merge_shards, merge_tensor, shard_tensor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _tool():
    spec = importlib.util.spec_from_file_location("convert_checkpoint", os.path.join(ROOT, "tools", "convert_checkpoint.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _shard_state(full: dict, cfg, tp: int, pp: int, ep: int) -> dict:
    """What every (tp_rank, pp_rank, ep_rank) would hold: slices of the unsharded state, stage-local layer numbers, the
    embedding on the first stage, final norm + head on the last."""
    L = cfg.num_hidden_layers
    out = {}
    for t in range(tp):
        for s in range(pp):
            ids = distribute_layers(L, pp, s)
            for e in range(ep):
                st = {}
                for name, v in full.items():
                    m = re.match(r"decoder_layers\.(\d+)\.(.*)", name)
                    if m:
                        g = int(m.group(1))
                        if g not in ids:
                            continue
                        local = f"decoder_layers.{ids.index(g)}.{m.group(2)}"
                    else:
                        if name == "embedding.weight" and s != 0 and not (cfg.tie_word_embeddings and s == pp - 1):
                            continue
                        if name in ("final_norm.weight", "final_proj.weight") and s != pp - 1:
                            continue
                        if name == "embedding.weight" and s == pp - 1 and s != 0:
                            continue                                      # the last stage keeps its tied copy as final_proj
                        local = name
                    st[local] = shard_tensor(name, v, cfg, tp, t, ep, e).clone()
                out[(t, s, e)] = st
    return out


CASES = [(Llama, "llama", dict(), 2, 1, 1), (Llama, "llama", dict(num_hidden_layers=5), 1, 3, 1),
         (Qwen3, "qwen3", dict(tie_word_embeddings=True, num_hidden_layers=4), 2, 2, 1),
         (Qwen3MoE, "qwen3_moe", dict(), 2, 1, 2), (Qwen3MoE, "qwen3_moe", dict(num_hidden_layers=3), 1, 2, 4)]


@pytest.mark.parametrize("cls,mt,kw,tp,pp,ep", CASES)
def test_merge_is_the_inverse_of_sharding(cls, mt, kw, tp, pp, ep):
    torch.manual_seed(0)
    cfg = TinyCfg(model_type=mt, **kw)
    full = {k: v.detach().clone() for k, v in cls(cfg, ParallelContext()).state_dict().items()}
    shards = _shard_state(full, cfg, tp, pp, ep)
    assert len(shards) == tp * pp * ep
    merged = merge_shards(shards, cfg, tp, pp, ep)
    assert set(merged) == set(full)
    for k in full:
        assert torch.equal(merged[k], full[k]), k


def test_merge_guards():
    cfg = TinyCfg(model_type="llama")
    full = Llama(cfg, ParallelContext()).state_dict()
    shards = _shard_state(full, cfg, 2, 1, 1)
    with pytest.raises(KeyError, match="missing shards"):
        merge_shards({k: v for k, v in shards.items() if k[0] == 0}, cfg, 2, 1, 1)
    shards[(1, 0, 0)]["final_norm.weight"] = shards[(1, 0, 0)]["final_norm.weight"] + 1        # replicated copies disagree
    with pytest.raises(ValueError, match="replicated"):
        merge_shards(shards, cfg, 2, 1, 1)
    with pytest.raises(ValueError, match="expected 2 tensor-parallel shards"):
        merge_tensor("embedding.weight", [torch.zeros(2, 2)], cfg, 2, 1)
    with pytest.raises(ValueError, match="expert-parallel shards"):
        merge_tensor("decoder_layers.0.moe.experts.down_proj", [torch.zeros(1, 2, 2)], cfg, 1, 2)


def _write_shards(tmp, shards, tp, pp, ep, step=7):
    os.makedirs(tmp, exist_ok=True)
    for (t, s, e), st in shards.items():
        name = f"weights_tp_rank_world_size={t}_{tp}_pp_rank_world_size={s}_{pp}" + (f"_ep_rank_world_size={e}_{ep}" if ep > 1 else "")
        torch.save({"model": st, "optimizer": None, "trained_steps": step, "trained_tokens": step * 100, "rng": None,
                    "layout": {"tp": tp, "pp": pp, "ep": ep}}, os.path.join(tmp, name + ".pth"))


@pytest.mark.parametrize("name,tp,pp,ep", [("tiny-qwen3", 2, 2, 1), ("tiny-qwen3-moe", 2, 1, 2)])
def test_tool_exports_hf_that_loads_back(tmp_path, name, tp, pp, ep):
    from scaletorch_b200.trainer import ScaleTorchArguments, create_model

    tool = _tool()
    cfg_args = ScaleTorchArguments(model_name_or_path=name, micro_batch_size=1, sequence_length=16, dtype="float32", use_cpu=True)
    torch.manual_seed(3)
    model, mc = create_model(cfg_args, torch.float32, torch.device("cpu"))
    full = {k: v.detach().clone() for k, v in model.state_dict().items()}
    _write_shards(str(tmp_path / "7"), _shard_state(full, mc, tp, pp, ep), tp, pp, ep)
    paths, layout = tool.find_shards(str(tmp_path / "7"))
    assert layout == (tp, pp, ep) and len(paths) == tp * pp * ep
    out = tmp_path / "export"
    assert tool.main(["--ckpt_dir", str(tmp_path / "7"), "--model", name, "--to", "hf", "--out", str(out)]) == 0
    assert (out / "model.safetensors").is_file() and (out / "config.json").is_file()
    cfg2 = ScaleTorchArguments(model_name_or_path=str(out), micro_batch_size=1, sequence_length=16, dtype="float32", use_cpu=True,
                               load_pretrained_weights=True, seed=99)
    torch.manual_seed(99)
    model2, _ = create_model(cfg2, torch.float32, torch.device("cpu"))
    ids = torch.randint(0, mc.vocab_size, (1, 16))
    assert rel_err(model2(ids), model(ids)) < 1e-6


def test_tool_writes_an_unsharded_training_checkpoint(tmp_path):
    from scaletorch_b200.trainer import ScaleTorchArguments, create_model
    from scaletorch_b200.utils.checkpoint import CheckpointManager

    tool = _tool()
    args = ScaleTorchArguments(model_name_or_path="tiny-llama", micro_batch_size=1, sequence_length=16, dtype="float32", use_cpu=True)
    torch.manual_seed(1)
    model, mc = create_model(args, torch.float32, torch.device("cpu"))
    full = {k: v.detach().clone() for k, v in model.state_dict().items()}
    _write_shards(str(tmp_path / "9"), _shard_state(full, mc, 2, 1, 1), 2, 1, 1, step=9)
    out = tmp_path / "merged"
    assert tool.main(["--ckpt_dir", str(tmp_path / "9"), "--model", "tiny-llama", "--to", "unsharded", "--out", str(out),
                      "--dtype", "float32"]) == 0
    torch.manual_seed(5)
    fresh, _ = create_model(ScaleTorchArguments(model_name_or_path="tiny-llama", micro_batch_size=1, sequence_length=16,
                                                dtype="float32", use_cpu=True, seed=5), torch.float32, torch.device("cpu"))
    ck = CheckpointManager()
    steps, tokens = ck.load_checkpoint(fresh, None, str(out))
    assert (steps, tokens) == (9, 900) and ck.last_extra == {"converted_from": {"tp": 2, "pp": 1, "ep": 1}}
    ids = torch.randint(0, mc.vocab_size, (1, 16))
    assert rel_err(fresh(ids), model(ids)) < 1e-6


def test_tool_rejects_empty_and_mixed_directories(tmp_path):
    tool = _tool()
    with pytest.raises(FileNotFoundError):
        tool.find_shards(str(tmp_path))
    for n in ("weights_tp_rank_world_size=0_2_pp_rank_world_size=0_1.pth", "weights_tp_rank_world_size=0_4_pp_rank_world_size=0_1.pth"):
        torch.save({}, tmp_path / n)
    with pytest.raises(ValueError, match="mixes layouts"):
        tool.find_shards(str(tmp_path))
    shutil.rmtree(tmp_path, ignore_errors=True)


def test_real_tp2_pp2_run_converts_and_continues_in_one_process(tmp_path):
    """tools/train.py on 4 gloo ranks (TP2 x PP2, tied embeddings) saves its shards; the converted checkpoint resumes
    in ONE process and the loss continues from where the parallel run left it."""
    from test_train_cli_cpu import _losses, _run

    tool = _tool()
    const_lr = ["--lr_scheduler_type", "step", "--step_size", "1000", "--log_lag", "0"]
    model = ["--model_name_or_path", "tiny-qwen3", "--num_attention_heads", "4", "--num_key_value_heads", "2"]
    work = tmp_path / "par"
    out = _run(model + const_lr + ["--tensor_parallel_size", "2", "--pipeline_parallel_size", "2", "--gradient_accumulation_steps", "2",
                                   "--total_train_steps", "6", "--work_dir", str(work), "--save_model_checkpoint", "true",
                                   "--save_frequency", "6"], nproc=4)
    par = _losses(out)
    assert len(par) == 6, out[-2000:]
    paths, layout = tool.find_shards(str(work / "6"))
    assert layout == (2, 2, 1) and len(paths) == 4
    merged = tmp_path / "merged"
    cfg_dir = tmp_path / "cfg"
    # the CLI overrides (4 heads, 2 kv heads) are part of the model: give the converter a config that carries them
    import json

    src = os.path.join(ROOT, "configs", "tiny-qwen3")
    shutil.copytree(src, cfg_dir)
    c = json.load(open(cfg_dir / "config.json"))
    c.update(num_attention_heads=4, num_key_value_heads=2)
    json.dump(c, open(cfg_dir / "config.json", "w"))
    assert tool.main(["--ckpt_dir", str(work / "6"), "--model", str(cfg_dir), "--to", "unsharded", "--out", str(merged)]) == 0
    out1 = _run(["--model_name_or_path", str(cfg_dir)] + const_lr + ["--gradient_accumulation_steps", "2", "--total_train_steps", "8",
                "--work_dir", str(tmp_path / "single"), "--resume_path", str(merged)])
    cont = _losses(out1)
    assert [s for s, _ in cont] == [7, 8], out1[-2000:]
    # same model, fresh optimiser state, and the data stream continues where the parallel run stopped (the converter
    # carries the position over in samples): the loss on the NEXT, unseen batches stays where the run had brought it.
    # (Restarting the data at sample 0 would show ~4.0 here: the tiny model has memorised the batches it trained on.)
    assert abs(cont[0][1] - par[-1][1]) < 0.5, (par, cont)
    # Without the converter: --resume_reshard reads the four shard files directly and re-cuts weights AND AdamW moments
    # (tied embedding on both pipeline ends, TP-sliced projections), so the single process continues the trajectory of
    # the TP2 x PP2 run itself.
    from test_train_cli_cpu import _run_many

    par_args = model + const_lr + ["--tensor_parallel_size", "2", "--pipeline_parallel_size", "2", "--gradient_accumulation_steps", "2",
                                   "--total_train_steps", "8", "--work_dir", str(tmp_path / "straight"), "--save_frequency", "0"]
    one_args = ["--model_name_or_path", str(cfg_dir)] + const_lr + ["--gradient_accumulation_steps", "2", "--total_train_steps", "8",
                "--work_dir", str(tmp_path / "direct"), "--save_frequency", "0", "--resume_path", str(work / "6"), "--resume_reshard", "true"]
    straight_out, direct_out = _run_many([(par_args, 4), (one_args, 1)])
    straight, direct = dict(_losses(straight_out)), dict(_losses(direct_out))
    assert "resharded the weights and AdamW moments" in direct_out and sorted(direct) == [7, 8]
    assert all(abs(direct[s] - straight[s]) < 2e-3 for s in (7, 8)), (direct, {s: straight[s] for s in (7, 8)})
