"""Continue a prompt with a model of this framework (sanity check of trained / exported weights).

    python tools/generate.py --model_name_or_path export/step300 --load_pretrained_weights true \
        --prompt "The capital of France is" --max_new_tokens 32 [--temperature 0.8 --top_k 50]
    python tools/generate.py --model_name_or_path tiny-qwen3 --prompt_ids 5,17,42 --max_new_tokens 8 --use_cpu true

Takes the trainer's flags (``ScaleTorchArguments``: device, dtype, tensor parallelism under torchrun, …) plus the five
generation flags above.  ``--prompt`` needs a tokenizer (``--tokenizer_name_or_path``, default: the model directory);
``--prompt_ids`` does not.  Greedy when ``--temperature`` is 0 (default).  Uses ``DecoderLM.generate``: the training
forward, no KV cache -- for checking weights, not for serving.  Prints one JSON line with the token ids (and the text).
"""
from __future__ import annotations

import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from transformers import HfArgumentParser  # noqa: E402

from scaletorch_b200.trainer import (ScaleTorchArguments, cleanup_distributed, create_model, get_dtype,  # noqa: E402
                                     initialize_distributed_training, validate_config)
from scaletorch_b200.utils.misc import set_all_seed  # noqa: E402


def _pop(argv: list[str], flag: str, default=None, cast=str):
    if flag in argv:
        i = argv.index(flag)
        val = cast(argv[i + 1])
        del argv[i:i + 2]
        return val
    return default


def main(argv=None) -> dict:
    argv = list(sys.argv[1:] if argv is None else argv)
    if "-h" in argv or "--help" in argv:
        print(__doc__)
        print("generation flags: --prompt TEXT | --prompt_ids 1,2,3   --max_new_tokens N (16)   --temperature T (0 = greedy)   --top_k K\n"
              "model / layout flags (the trainer's):\n")
        HfArgumentParser(ScaleTorchArguments).parse_args_into_dataclasses(args=["--help"])
    prompt = _pop(argv, "--prompt")
    prompt_ids = _pop(argv, "--prompt_ids")
    max_new = _pop(argv, "--max_new_tokens", 16, int)
    temperature = _pop(argv, "--temperature", 0.0, float)
    top_k = _pop(argv, "--top_k", None, int)
    if (prompt is None) == (prompt_ids is None):
        raise SystemExit("tools/generate.py: give exactly one of --prompt / --prompt_ids")
    (config,) = HfArgumentParser(ScaleTorchArguments).parse_args_into_dataclasses(args=argv)
    if config.pipeline_parallel_size > 1 or config.context_parallel_size > 1:
        raise SystemExit("tools/generate.py: use pipeline_parallel_size = context_parallel_size = 1")
    dtype = get_dtype(config)
    _, rank, world, device = initialize_distributed_training(config)
    validate_config(config, world)
    try:
        set_all_seed(config.seed)
        model, mc = create_model(config, dtype, device)
        raw = model.module if hasattr(model, "module") else model
        tok = None
        if prompt is not None:
            from transformers import AutoTokenizer

            name = config.tokenizer_name_or_path if config.tokenizer_name_or_path != "facebook/opt-125m" else config.model_name_or_path
            tok = AutoTokenizer.from_pretrained(name)
            ids = tok(prompt, return_tensors="pt")["input_ids"]
        else:
            ids = torch.tensor([[int(t) for t in prompt_ids.split(",") if t.strip()]], dtype=torch.long)
        if int(ids.max()) >= mc.vocab_size or int(ids.min()) < 0:
            raise SystemExit(f"prompt ids must lie in [0, {mc.vocab_size})")
        gen = torch.Generator(device="cpu").manual_seed(config.seed)
        out = raw.generate(ids.to(device), max_new, temperature=temperature, top_k=top_k,
                           eos_token_id=getattr(mc, "eos_token_id", None) if tok is not None else None,
                           generator=gen if device.type == "cpu" else None)
        res = {"prompt_ids": ids[0].tolist(), "new_ids": out[0, ids.shape[1]:].tolist()}
        if tok is not None:
            res["text"] = tok.decode(out[0].tolist())
        if rank == 0:
            print(json.dumps(res), flush=True)
        return res
    finally:
        cleanup_distributed()


if __name__ == "__main__":
    main()
