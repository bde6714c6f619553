"""Reference arm of bench.py: runs the UNMODIFIED jianzhnie/ScaleTorch installed under ``baseline/_ref`` through
its own public API (``ScaleTorchArguments`` → ``initialize_distributed_training`` → ``create_model`` →
``create_optimizer`` → ``train_step`` → ``clip_gradients`` → ``optimizer.step`` → ``model.reset``), i.e. the
sequence of its ``tools/train.py:_run_training_loop``, on the same config and metric as our arm.

Only the *data source* is ours: the reference's ``MicroBatchDataLoader`` needs a HF dataset + tokenizer
download (no network here), so a synthetic-token iterator with the loader's batch format feeds
``train_step`` -- the same substitution the reference's own ``tools/bench_single.py:63`` makes.
None of our models, kernels or engine are imported on this path.
"""
from __future__ import annotations

import json
import os
import sys


def random_checkpoint(model_dir: str, local_rank: int, limit_gb: float = 2.6):
    """A random-weight HF-layout ``model.safetensors`` next to a copy of ``config.json`` (under /tmp), for models with
    tied embeddings only.  Why: without a checkpoint the reference materialises with ``to_empty`` + ``reset_parameters``
    (`scaletorch/utils/checkpoint.py:90-97`), which unties the LM head and never initialises it (`model_qwen3.py:313-322`
    skips ``final_proj`` when ``tie_word_embeddings``): 751.6 M instead of 596.0 M parameters for Qwen3-0.6B and a NaN
    loss from the uninitialised matrix.  Its stock checkpoint path (`checkpoint.py:99-142`) ties the head
    (``_handle_final_projection``) and then re-randomises everything with ``reset_parameters`` -- so a file of random
    weights gives exactly "random-init weights of that architecture" through the reference's own loader.
    Returns (dir to pass as ``model_name_or_path``, note)."""
    import shutil

    import torch

    cfg = json.load(open(os.path.join(model_dir, "config.json")))
    if not cfg.get("tie_word_embeddings", False) or cfg.get("model_type") not in ("qwen3", "llama", "qwen2"):
        return model_dir, "no-checkpoint init (untied head: nothing to work around)"
    H, L, V = cfg["hidden_size"], cfg["num_hidden_layers"], cfg["vocab_size"]
    nh, nkv, I = cfg["num_attention_heads"], cfg["num_key_value_heads"], cfg["intermediate_size"]
    hd = cfg.get("head_dim") or H // nh
    per_layer = {"input_layernorm": (H,), "post_attention_layernorm": (H,), "mlp.gate_proj": (I, H), "mlp.up_proj": (I, H),
                 "mlp.down_proj": (H, I), "self_attn.q_proj": (nh * hd, H), "self_attn.k_proj": (nkv * hd, H),
                 "self_attn.v_proj": (nkv * hd, H), "self_attn.o_proj": (H, nh * hd)}
    if cfg["model_type"] in ("qwen3", "qwen2"):
        per_layer.update({"self_attn.q_norm": (hd,), "self_attn.k_norm": (hd,)})
    n = V * H + H + L * sum(int(torch.Size(s).numel()) for s in per_layer.values())
    if n * 2 / 1e9 > limit_gb:
        return model_dir, f"no-checkpoint init (random checkpoint would be {n * 2 / 1e9:.1f} GB)"
    out = os.path.join("/tmp", "stb_ref_ckpt", os.path.basename(model_dir.rstrip("/")))
    done = os.path.join(out, ".done")
    if local_rank == 0 and not os.path.exists(done):
        from safetensors.torch import save_file

        os.makedirs(out, exist_ok=True)
        for f in os.listdir(model_dir):
            if f.endswith(".json") or f.endswith(".txt"):
                shutil.copy(os.path.join(model_dir, f), os.path.join(out, f))
        g = torch.Generator().manual_seed(1234)
        mk = lambda shape: (torch.ones(shape) if len(shape) == 1 else torch.randn(shape, generator=g) * 0.02).to(torch.bfloat16)  # noqa: E731
        sd = {"model.embed_tokens.weight": mk((V, H)), "model.norm.weight": mk((H,))}
        for i in range(L):
            for k, shape in per_layer.items():
                sd[f"model.layers.{i}.{k}.weight"] = mk(shape)
        save_file(sd, os.path.join(out, "model.safetensors"))
        open(done, "w").write("ok")
    elif local_rank != 0:
        import time

        t0 = time.time()                      # the other ranks of the node wait for local rank 0's files
        while not os.path.exists(done) and time.time() - t0 < 300:
            time.sleep(0.2)
    return out, "random-weight checkpoint through the reference's HF loader (keeps the LM head tied)"


def main(args, root, ref_dir, ClockSampler, timed_region, build_config_dict, dist_env, baseline_table):
    sys.path.insert(0, ref_dir)
    os.environ.setdefault("FLASH_ATTEN", "1")
    os.environ.setdefault("DTYPE", "bfloat16")
    if args.cp > 1:
        os.environ.setdefault("CONTEXT_PARALLEL", "1")  # otherwise the reference silently skips the ring
    if args.sequence_parallel:
        os.environ.setdefault("SEQUENCE_PARALLEL", "1")
    import torch
    import torch.distributed as dist

    from scaletorch.parallel.process_group import process_group_manager as pgm
    from scaletorch.trainer import (ScaleTorchArguments, clip_gradients, create_model, create_optimizer, get_dtype,
                                    initialize_distributed_training, train_step)
    from scaletorch.utils.misc import get_num_params

    assert "scaletorch_b200" not in sys.modules
    rank, local_rank, world = dist_env()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    dp, cfgd = build_config_dict(args, world)
    model_dir, init_note = random_checkpoint(os.path.join(root, "configs", args.model), local_rank)
    cfg = ScaleTorchArguments(
        model_name_or_path=model_dir, micro_batch_size=args.micro_batch_size, sequence_length=args.seq_len,
        gradient_accumulation_steps=args.grad_accum, tensor_parallel_size=args.tp, context_parallel_size=args.cp,
        pipeline_parallel_size=args.pp, expert_parallel_size=args.ep, data_parallel_size=dp,
        gradient_checkpointing=args.gradient_checkpointing, learning_rate=3e-4, max_grad_norm=1.0, seed=1)
    dtype = get_dtype(cfg)
    _, _, _, _, device = initialize_distributed_training(cfg)
    if device.type == "cuda":
        torch.cuda.set_device(device)
    if world > 1:
        dist.barrier()  # local rank 0 has finished writing the checkpoint
    model, model_config = create_model(cfg, dtype, device)
    model.train()
    optimizer = create_optimizer(model, cfg, device)
    num_params = get_num_params(model)
    s_local = args.seq_len // args.cp
    gen = torch.Generator().manual_seed(1 + rank)
    vocab = model_config.vocab_size

    class Synthetic:
        """Batches in the reference loader's format; ``pinned=True`` keeps them in pinned host memory."""

        gradient_accumulation_steps = args.grad_accum

        def __init__(self, n, on_device):
            self.b, self.i = [], 0
            for _ in range(n):
                ids = torch.randint(0, vocab, (args.micro_batch_size, s_local + 1), generator=gen)
                x, y = ids[:, :-1].contiguous(), ids[:, 1:].contiguous()
                if on_device:
                    x, y = x.to(device), y.to(device)
                else:
                    x, y = x.pin_memory(), y.pin_memory()
                self.b.append({"input_ids": x, "target_ids": y, "position_ids": None, "hidden_states": None})

        def __next__(self):
            x = self.b[self.i % len(self.b)]
            self.i += 1
            return x

    def step_with(src):
        optimizer.zero_grad(set_to_none=True)
        loss = train_step(model=model, data_loader=src, device=device, dtype=dtype, scaler=None,
                          gradient_accumulation_steps=args.grad_accum,
                          gradient_checkpointing=args.gradient_checkpointing)
        clip_gradients(model, 1.0)
        optimizer.step()
        if hasattr(model, "reset"):
            model.reset()
        return loss

    staged = Synthetic((args.steps + args.warmup) * args.grad_accum, on_device=True)
    for _ in range(args.warmup):
        step_with(staged)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms = timed_region(torch, dist, world, lambda: step_with(staged), args.steps)
    clocks = sampler.stop() if rank == 0 else {}
    e2e = None
    if not args.skip_e2e:
        host = Synthetic((args.steps + 1) * args.grad_accum, on_device=False)
        losses = []
        step = lambda: losses.append(float(step_with(host)))  # noqa: E731  (reference train_step already .item()s)
        step()
        ms_e2e = timed_region(torch, dist, world, step, args.steps)
        tokens = cfgd["global_batch"] * args.seq_len
        per_rank_tokens = args.micro_batch_size * args.grad_accum * s_local
        e2e = {"value": tokens * args.steps / (ms_e2e / 1e3), "unit": "tokens/s", "ms_per_step": ms_e2e / args.steps,
               "h2d_bytes_per_step": per_rank_tokens * 8 * 2, "d2h_bytes_per_step": 4 * args.grad_accum,
               "last_loss": losses[-1]}
    if rank == 0:
        tokens = cfgd["global_batch"] * args.seq_len
        value = tokens * args.steps / (ms / 1e3)
        base = baseline_table.get(args.model)
        print(json.dumps({
            "metric": "tokens/s (whole job; tokens/s/GPU = value / n_gpus) + MFU", "value": value, "unit": "tokens/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": (value / (base * world)) if base else None,
            "dtype": str(dtype).replace("torch.", ""), "data": "synthetic tokens, random-init weights",
            "impl": "reference", "config": cfgd, "tokens_per_s_per_gpu": value / world, "num_params": num_params,
            "clocks": clocks, "e2e": e2e, "gpu_launches": 0,
            "init": init_note,
            "notes": "unmodified reference (PyTorch + cuBLAS + SDPA + NCCL); colorama replaced by a stand-in module",
        }), flush=True)
    if world > 1 and dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()
