"""Held-out loss / perplexity of a model under any data-parallel width.

    python tools/evaluate.py --model_name_or_path export/step300 --load_pretrained_weights true --token_file val.npy \
        --sequence_length 2048 --micro_batch_size 4 [--max_batches 100]
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 tools/evaluate.py … --data_parallel_size 8

Takes the trainer's flags (``ScaleTorchArguments``) plus ``--max_batches``.  Data: ``--token_file`` (flat int32 token ids,
``.npy`` or raw ``.bin``; windows of ``S + 1`` tokens, no overlap), a dataset through ``--dataset_name`` /
``--tokenizer_name_or_path``, or ``--synthetic_data true``.  Every data-parallel rank evaluates its shard of the windows;
token counts and summed losses meet in ONE ``all_reduce_dict`` at the end (``scaletorch_b200.dist``), rank 0 prints one
JSON line: ``{"loss", "perplexity", "tokens", "batches", "tokens_per_s"}``.  Tensor / context parallel layouts work as in
training (the loss is already reduced over them); the stand-alone tool does not take pipeline
layouts (the trainer's ``--eval_interval`` does).
"""
from __future__ import annotations

import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from transformers import HfArgumentParser  # noqa: E402

from scaletorch_b200.data.dataloader import TokenWindows, load_tokens  # noqa: E402
from scaletorch_b200.trainer import (ScaleTorchArguments, cleanup_distributed, create_model, get_dtype,  # noqa: E402
                                     initialize_distributed_training, validate_config)
from scaletorch_b200.trainer.evaluation import evaluate_loss, reduce_eval_stats  # noqa: E402


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
        res = reduce_eval_stats(evaluate_loss(model, loader, device, max_batches), device)
        out = {"loss": res["loss"], "perplexity": res["perplexity"], "tokens": res["tokens"], "batches": res["batches"],
               "tokens_per_s": res["tokens"] / max(res["seconds"], 1e-9), "model": config.model_name_or_path, "world_size": world}
        if rank == 0:
            print(json.dumps(out), flush=True)
        return out
    finally:
        cleanup_distributed()


if __name__ == "__main__":
    main()
