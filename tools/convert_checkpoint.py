"""Merge the per-(tp, pp, ep) training checkpoints of one step into ONE model and export it.

    python tools/convert_checkpoint.py --ckpt_dir work_dir/300 --model Qwen3-8B --to hf --out export/qwen3-8b-step300
    python tools/convert_checkpoint.py --ckpt_dir work_dir/300 --model Qwen3-8B --to unsharded --out work_dir/300_merged

``CheckpointManager`` writes one ``weights_tp_rank_world_size=…_pp_rank_world_size=…[_ep_rank_world_size=…].pth`` per
coordinate and resumes only under the same layout (as the reference does, ``scaletorch/utils/checkpoint.py:467-560``).
This tool is the way out of a layout: it reads every shard of a step, undoes the tensor-parallel slicing, the pipeline
stage numbering and the expert sharding (``utils/sharding.py::merge_shards``), and writes either

* ``--to hf``: ``model.safetensors`` (+ ``config.json``) in the HuggingFace layout -- loadable by ``transformers``, by
  this framework under ANY layout (``--load_pretrained_weights true``), and by serving stacks;
* ``--to unsharded``: a training checkpoint for tp = pp = ep = 1 (model weights, step and token counters; optimiser
  state is NOT converted -- it restarts, as after any fine-tuning hand-over; the position in the data stream IS carried
  over, in samples, so the continued run does not see the same batches again).

Pure CPU; needs no process group and no GPU.
"""
from __future__ import annotations

import argparse
import json
import os
import shutil
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from scaletorch_b200.trainer.config import resolve_model_config  # noqa: E402
from scaletorch_b200.utils.checkpoint import hf_state_from_unsharded  # noqa: E402
from scaletorch_b200.utils.checkpoint import find_shards, merge_checkpoint  # noqa: E402,F401

def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--ckpt_dir", required=True, help="directory of ONE step: work_dir/<step>")
    ap.add_argument("--model", required=True, help="model name or directory (config.json), as given to --model_name_or_path")
    ap.add_argument("--to", choices=["hf", "unsharded"], default="hf")
    ap.add_argument("--out", required=True)
    ap.add_argument("--dtype", default=None, choices=[None, "bfloat16", "float32", "float16"], help="cast weights on export")
    a = ap.parse_args(argv)
    cfg = resolve_model_config(a.model)
    full, meta = merge_checkpoint(a.ckpt_dir, cfg)
    if a.dtype:
        dt = getattr(torch, a.dtype)
        full = {k: (v.to(dt) if v.is_floating_point() else v) for k, v in full.items()}
    os.makedirs(a.out, exist_ok=True)
    n = sum(v.numel() for k, v in full.items() if not (k == "final_proj.weight" and getattr(cfg, "tie_word_embeddings", False)))
    if a.to == "hf":
        from safetensors.torch import save_file

        sd = hf_state_from_unsharded(full, cfg)
        save_file(sd, os.path.join(a.out, "model.safetensors"), metadata={"format": "pt", **{k: json.dumps(v) for k, v in meta.items()}})  # (strings only)
        src_cfg = next((os.path.join(d, "config.json") for d in (a.model, os.path.join(os.path.dirname(os.path.dirname(
            os.path.abspath(__file__))), "configs", a.model)) if os.path.isfile(os.path.join(d, "config.json"))), None)
        if src_cfg:
            shutil.copy(src_cfg, os.path.join(a.out, "config.json"))
        print(f"wrote {len(sd)} tensors ({n / 1e6:.1f} M parameters) from layout {meta['source_layout']} to {a.out}/model.safetensors")
    else:
        payload = {"model": full, "optimizer": None, "trained_steps": meta["trained_steps"],
                   "trained_tokens": meta["trained_tokens"], "rng": None, "layout": {"tp": 1, "pp": 1, "ep": 1},
                   "extra": {"converted_from": meta["source_layout"],
                             **({"data_loader": meta["data_loader"]} if "data_loader" in meta else {})}}
        path = os.path.join(a.out, "weights_tp_rank_world_size=0_1_pp_rank_world_size=0_1.pth")
        torch.save(payload, path)
        print(f"wrote {path} ({n / 1e6:.1f} M parameters, step {meta['trained_steps']})")
    return 0


if __name__ == "__main__":
    sys.exit(main())
