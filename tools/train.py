#!/usr/bin/env python
"""Training entry point.

    torchrun --nproc_per_node 8 --master-addr 127.0.0.1 tools/train.py \
        --model_name_or_path Qwen3-8B --tensor_parallel_size 2 --context_parallel_size 2 --data_parallel_size 2 \
        --sequence_parallel True --micro_batch_size 1 --sequence_length 4096 --synthetic_data True \
        --total_train_steps 20

Same flags and console output as the reference trainer (``tools/train.py:55-453``: config → dist init →
data loader → model → optimiser / scheduler → loop with clip, step, log, checkpoint → cleanup with
``performance_logs_{rank}_{ts}.json``).  Differences: device-timed steps (CUDA events), one host sync per
step (the loss), global-norm clipping folded into the fused optimiser, RNG / data position in checkpoints.
"""
from __future__ import annotations

import gc
from collections import deque
import os
import sys
import time
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from transformers import HfArgumentParser  # noqa: E402

from scaletorch_b200.data import MicroBatchDataLoader  # noqa: E402
from scaletorch_b200.parallel.process_group import process_group_manager as pgm  # noqa: E402
from scaletorch_b200.trainer import (AsyncScalarReader, ScaleTorchArguments, cleanup_distributed, clip_gradients, create_lr_scheduler,  # noqa: E402
                                     create_model, create_optimizer, get_dtype, get_tensor_shapes,
                                     initialize_distributed_training, log_training_metrics, training_step,
                                     validate_config)
from scaletorch_b200.trainer.metrics import is_logging_rank  # noqa: E402
from scaletorch_b200.trainer.optimizer import FusedAdamW  # noqa: E402
from scaletorch_b200.utils.checkpoint import CheckpointManager  # noqa: E402
from scaletorch_b200.utils.device import empty_cache, is_accelerator_available  # noqa: E402
from scaletorch_b200.utils.logger_utils import get_logger  # noqa: E402
from scaletorch_b200.utils.misc import (average_loss_across_dp_cp_ranks, get_num_params, rank_print, set_all_seed,  # noqa: E402
                                        to_readable_format)
from scaletorch_b200.utils.monitor import PerformanceMonitor  # noqa: E402

logger = get_logger("scaletorch_b200.train")



def _import_wandb():
    """wandb is imported only when ``--use_wandb true`` asks for it: the import alone costs seconds on every rank."""
    try:
        import wandb
    except Exception:  # pragma: no cover - optional dependency
        rank_print("--use_wandb is set but wandb cannot be imported; continuing without it")
        return None
    return wandb


def report_dropped_pairs(model, step: int) -> None:
    """The peer-memory EP path has a static receive capacity (``--ep_capacity_factor`` / SCALETORCH_B200_EP_CAPACITY);
    pairs beyond it are dropped on the sender.  Make that visible: every 50 steps each rank that dropped something says so
    (one host read per MoE layer, off the critical path of the lagged logger) and the counters are reset."""
    raw = model
    while hasattr(raw, "module") and isinstance(raw.module, torch.nn.Module):
        raw = raw.module
    total = 0
    for m in raw.modules():
        d = getattr(m, "dropped_pairs", None)
        if isinstance(d, torch.Tensor):
            total += int(d.item())
            m.dropped_pairs = None
    if total > 0:
        rank_print(f"[step {step}] expert-parallel capacity exceeded: {total} (token, expert) pairs dropped since the last "
                   f"report; raise --ep_capacity_factor", is_print_rank=True)


class StopOnSignal:
    """Graceful preemption (``--save_on_signal true``): SIGTERM / SIGUSR1 only set a flag; every
    ``--signal_check_interval`` steps the ranks take the maximum of their flags over a small gloo group (signals reach
    ranks at different times, and all of them have to stop after the SAME step because the step is full of collectives);
    when the maximum is set, the caller writes a checkpoint and leaves the loop.  The reference has no signal handling
    (SURVEY §5.3)."""

    def __init__(self, enabled: bool, interval: int, world_size: int):
        self.enabled, self.interval = bool(enabled), max(1, int(interval))
        self.requested = False
        self.group = None
        if not self.enabled:
            return
        import signal

        for sig in (signal.SIGTERM, signal.SIGUSR1):
            signal.signal(sig, self._on_signal)
        if world_size > 1:
            self.group = dist.new_group(backend="gloo")          # every rank; host-side, independent of the NCCL stream

    def _on_signal(self, signum, frame) -> None:
        self.requested = True

    def should_stop(self, step: int) -> bool:
        if not self.enabled or (self.group is not None and step % self.interval != 0):
            return False
        if self.group is None:
            return self.requested
        flag = torch.tensor([int(self.requested)], dtype=torch.int64)
        dist.all_reduce(flag, op=dist.ReduceOp.MAX, group=self.group)
        return bool(flag.item())


def build_native_loader(config, model_config):
    """C++ micro-batch producer over a flat token array: a ``--token_file`` (int32 ``.npy`` or raw ``.bin``) or, with
    ``--synthetic_data``, random ids.  Same batch dict and ``global_batch_size`` attribute as ``MicroBatchDataLoader``."""
    import numpy as np

    from scaletorch_b200.data import NativeTokenLoader
    from scaletorch_b200.parallel.context_parallel import zigzag

    S = config.sequence_length
    if config.token_file:
        arr = np.load(config.token_file) if config.token_file.endswith(".npy") else np.fromfile(config.token_file, dtype=np.int32)
        tokens = torch.from_numpy(np.ascontiguousarray(arr, dtype=np.int32))
    else:
        g = torch.Generator().manual_seed(config.seed)
        n = (config.num_samples or 4096) * (S + 1)
        tokens = torch.randint(0, model_config.vocab_size, (n,), generator=g, dtype=torch.int32)
    data_world = pgm.data_world_size if pgm else 1
    loader = NativeTokenLoader(
        tokens, S, config.micro_batch_size, cp_rank=pgm.cp_rank if pgm else 0, cp_size=pgm.cp_world_size if pgm else 1,
        zigzag=zigzag.enabled(), data_rank=pgm.data_rank if pgm else 0, data_world=data_world, shuffle=True,
        seed=config.seed, gradient_accumulation_steps=config.gradient_accumulation_steps)
    loader.global_batch_size = config.micro_batch_size * config.gradient_accumulation_steps * data_world
    loader.state_dict = lambda: {"batches_served": loader.batches_served}
    return loader


def build_data_loader(config, model_config, device):
    if getattr(config, "native_loader", False):
        return build_native_loader(config, model_config)
    return MicroBatchDataLoader(
        micro_batch_size=config.micro_batch_size, sequence_length=config.sequence_length,
        dataset_name=config.dataset_name, tokenizer_name_or_path=config.tokenizer_name_or_path
        if config.tokenizer_name_or_path != "facebook/opt-125m" else config.model_name_or_path,
        gradient_accumulation_steps=config.gradient_accumulation_steps, device=device,
        num_workers=config.num_workers if not config.synthetic_data else 0, num_proc=config.num_proc,
        num_samples=config.num_samples, subset_name=config.subset_name, split=config.split,
        pin_memory=config.pin_memory, synthetic=config.synthetic_data, vocab_size=model_config.vocab_size,
        seed=config.seed)


def build_eval_loader(config, model_config, device):
    """The held-out loader of ``--eval_interval``: ``--eval_token_file`` windows, else the dataset's ``--eval_split``, else
    (synthetic data) sequences drawn with another seed than the training stream.  Never shuffled; ``None`` when evaluation
    is off."""
    if getattr(config, "eval_interval", 0) <= 0:
        return None
    dataset = None
    if config.eval_token_file:
        from scaletorch_b200.data.dataloader import TokenWindows, load_tokens

        dataset = TokenWindows(load_tokens(config.eval_token_file), config.sequence_length)
    synthetic = dataset is None and (config.synthetic_data or getattr(config, "native_loader", False) and not config.dataset_name)
    return MicroBatchDataLoader(
        micro_batch_size=config.micro_batch_size, sequence_length=config.sequence_length,
        dataset_name=config.dataset_name, tokenizer_name_or_path=config.tokenizer_name_or_path
        if config.tokenizer_name_or_path != "facebook/opt-125m" else config.model_name_or_path,
        device=device, num_workers=0, num_proc=config.num_proc, subset_name=config.subset_name, split=config.eval_split,
        pin_memory=config.pin_memory, synthetic=synthetic, vocab_size=model_config.vocab_size, seed=config.seed + 7919,
        num_samples=None if not synthetic else max(4096, 8 * config.eval_batches * config.micro_batch_size),
        dataset=dataset, shuffle=False)


def main(argv=None) -> None:
    world_size, monitor, run = 1, None, None
    try:
        (config,) = HfArgumentParser(ScaleTorchArguments).parse_args_into_dataclasses(args=argv)
        dtype = get_dtype(config)
        local_rank, global_rank, world_size, device = initialize_distributed_training(config)
        validate_config(config, world_size)
        log_rank = is_logging_rank()
        set_all_seed(config.seed)
        logger.info("dtype=%s device=%s world_size=%d layout=%s", dtype, device, world_size, pgm if pgm else "single")

        t0 = time.time()
        model, model_config = create_model(config, dtype, device)
        model.train()
        data_loader = build_data_loader(config, model_config, device)
        if world_size > 1:
            dist.barrier()
        tokens_per_step = data_loader.global_batch_size * config.sequence_length
        rank_print(f"init time: {time.time() - t0:.2f}s | Tokens per step: {to_readable_format(tokens_per_step)}",
                   is_print_rank=log_rank)
        wandb = _import_wandb() if (log_rank and config.use_wandb) else None
        if wandb is not None:
            run = wandb.init(project=config.project_name, name=f"{config.experiment_name}_{to_readable_format(tokens_per_step)}",
                             config={k: getattr(config, k) for k in ("tensor_parallel_size", "context_parallel_size",
                                     "pipeline_parallel_size", "data_parallel_size", "expert_parallel_size",
                                     "model_name_or_path", "learning_rate", "seed", "micro_batch_size",
                                     "gradient_accumulation_steps", "sequence_length")})
        num_params = get_num_params(model)
        rank_print(f"Number of parameters: {to_readable_format(num_params)}", is_print_rank=log_rank)
        optimizer = create_optimizer(model, config, device)
        total_steps = config.total_train_steps
        if total_steps is None and config.max_tokens is not None and tokens_per_step > 0:
            total_steps = config.max_tokens // tokens_per_step
        lr_scheduler = create_lr_scheduler(optimizer, config, total_steps)
        ckpt = CheckpointManager(async_save=getattr(config, "async_checkpoint", False))
        monitor = PerformanceMonitor(config=config, log_dir=config.work_dir)
        trained_tokens, step = 0, 0
        from scaletorch_b200.utils.checkpoint import prune_checkpoints, resolve_resume_path

        config.resume_path = resolve_resume_path(config.resume_path, config.work_dir)   # 'latest' -> newest complete one
        if config.resume_path and not ckpt.layout_matches(config.resume_path):
            if not getattr(config, "resume_reshard", False):
                raise SystemExit(f"{config.resume_path} was written under another TP / PP / EP layout; pass --resume_reshard true to "
                                 "merge and re-cut its weights for this layout (optimiser moments restart from zero)")
            step, trained_tokens = ckpt.load_resharded(model, config.resume_path, model_config, optimizer)
            if ckpt.moments_resharded:
                rank_print(f"resharded the weights and AdamW moments of {config.resume_path} from {ckpt.last_source_layout} to this "
                           "layout", is_print_rank=log_rank)
            else:
                rank_print(f"WARNING: resharded the weights of {config.resume_path} from {ckpt.last_source_layout} to this layout; "
                           "the AdamW moments restart from zero", is_print_rank=log_rank)
                if hasattr(optimizer, "step_count"):
                    optimizer.step_count = 0               # fresh moments need fresh bias corrections
        elif config.resume_path:
            step, trained_tokens = ckpt.load_checkpoint(model, optimizer, config.resume_path)
        if config.resume_path:
            sched_path = os.path.join(config.resume_path, "scheduler.pt")
            if os.path.exists(sched_path):
                lr_scheduler.load_state_dict(torch.load(sched_path, weights_only=False))
            dl_state = (getattr(ckpt, "last_extra", None) or {}).get("data_loader")
            if dl_state and hasattr(data_loader, "load_state_dict"):
                data_loader.load_state_dict(dl_state)
            rank_print(f"Loaded checkpoint at step {step}, trained_tokens={trained_tokens}", is_print_rank=log_rank)
        if world_size > 1:
            dist.barrier()
        tensor_shapes = get_tensor_shapes(config, model_config) if (pgm and pgm.pp_world_size > 1) else None

        from scaletorch_b200.trainer import GraphedTrainStep
        from scaletorch_b200.utils.tracing import ProfilerWindow, enable_nvtx, nvtx_range
        from scaletorch_b200.utils.watchdog import StepWatchdog, maybe_inject_fault

        if config.nvtx:
            enable_nvtx()
        prof = ProfilerWindow(config.profile_start_step, config.profile_end_step, config.trace_dir, global_rank)
        watchdog = StepWatchdog(config.step_timeout_s, config.work_dir or ".", global_rank).start()
        graphed = None
        if config.use_cuda_graph and GraphedTrainStep.supported(config, optimizer, model):
            graphed = GraphedTrainStep(model, optimizer, config, device)
            rank_print("whole-step CUDA graph replay enabled", is_print_rank=log_rank)
        from scaletorch_b200.trainer.evaluation import PeriodicEvaluator, format_eval_line

        def report_eval(st, result):
            rank_print(format_eval_line(st, result), is_print_rank=log_rank)
            if run is not None:
                run.log({"eval_loss": result["loss"], "eval_perplexity": result["perplexity"]}, step=st)

        evaluator = PeriodicEvaluator(model, build_eval_loader(config, model_config, device), device,
                                      getattr(config, "eval_interval", 0), getattr(config, "eval_batches", 8), report_eval,
                                      tensor_shapes=tensor_shapes, dtype=dtype)
        stopper = StopOnSignal(getattr(config, "save_on_signal", False), getattr(config, "signal_check_interval", 10), world_size)
        monitor.start()
        max_steps = config.total_train_steps or float("inf")
        log_lag = max(0, int(getattr(config, "log_lag", 1)))
        pending: deque = deque()
        loss_reader, norm_reader = AsyncScalarReader(log_lag + 2), AsyncScalarReader(log_lag + 2)

        def flush_one():
            st, lr_, gn, handle, tokens_ = pending.popleft()
            stats = monitor.resolve_iteration(handle)
            loss = loss_reader.pop()
            gn = norm_reader.pop() if gn == "reader" else gn
            if loss != loss or loss in (float("inf"), float("-inf")):
                # ranks that log a loss read the same averaged value and take the same branch; with pipeline stages only the
                # last one sees it, and the launcher ends the other ranks when that one exits
                msg = f"non-finite loss ({loss}) at step {st}"
                if getattr(config, "nonfinite_loss", "abort") == "abort":
                    raise FloatingPointError(msg + ": aborting (--nonfinite_loss warn continues instead); resume from the "
                                             "last checkpoint with a lower learning rate or inspect the data of this step")
                rank_print("WARNING: " + msg, is_print_rank=log_rank)
            log_training_metrics(config, model_config, st, loss, lr_, gn, stats["iteration_time"], tokens_per_step, tokens_,
                                 num_params, world_size, logger=logger, wandb_run=run)
            if st % 50 == 0:
                report_dropped_pairs(model, st)

        while config.max_tokens is None or trained_tokens < config.max_tokens:
            prof.step(step)
            maybe_inject_fault(global_rank, step)
            monitor.start_iteration(tokens_processed=tokens_per_step)
            with nvtx_range(f"step {step}"):
                if graphed is not None:
                    loss_t, grad_norm = graphed(data_loader), None
                else:
                    optimizer.zero_grad(set_to_none=True)
                    if hasattr(model, "reset"):
                        model.reset()
                    with nvtx_range("fwd+bwd"):
                        loss_t = training_step(model, data_loader, device, dtype, config, tensor_shapes)
                    grad_norm = clip_gradients(model, config.max_grad_norm, optimizer)
                    with nvtx_range("optimizer"):
                        optimizer.step()
            lr_scheduler.step()
            # lagged logging: step i is resolved (events, loss, grad norm read back) after step i+1 has been enqueued,
            # so the host never drains the device queue; --log_lag 0 restores fully synchronous logging
            handle = monitor.end_iteration_deferred()
            loss_reader.push(average_loss_across_dp_cp_ranks(loss_t, device, as_tensor=True))
            if isinstance(optimizer, FusedAdamW) and config.max_grad_norm and optimizer.last_grad_norm is not None:
                norm_reader.push(optimizer.last_grad_norm)
                grad_norm = "reader"
            trained_tokens += tokens_per_step
            step += 1
            pending.append((step, lr_scheduler.get_last_lr()[0], grad_norm, handle, trained_tokens))
            if step % 100 == 0 and is_accelerator_available():
                empty_cache()
                if step % 1000 == 0:
                    gc.collect()
            while len(pending) > log_lag:
                flush_one()
            watchdog.beat(step)
            if evaluator.due(step):
                while pending:                            # the training line of this step goes out before the eval line
                    flush_one()
                evaluator.run(step)
                watchdog.beat(step)
            if config.save_model_checkpoint and config.save_frequency > 0 and step % config.save_frequency == 0:
                while pending:                            # know this step's loss before publishing its weights (non-finite -> abort)
                    flush_one()
                out_dir = Path(config.work_dir) / f"{step}"
                ckpt.save_checkpoint(model, optimizer, step, trained_tokens, str(out_dir),
                                     extra={"data_loader": data_loader.state_dict()})
                if log_rank:
                    torch.save(lr_scheduler.state_dict(), out_dir / "scheduler.pt")
                if global_rank == 0 and getattr(config, "keep_last_checkpoints", 0) > 0:
                    # older COMPLETE checkpoints only; the one being written (other ranks may still be at it) is protected
                    for gone in prune_checkpoints(config.work_dir, config.keep_last_checkpoints, protect=str(out_dir)):
                        logger.info("removed old checkpoint %s", gone)
            if step >= max_steps:
                break
            if stopper.should_stop(step):
                while pending:
                    flush_one()
                out_dir = Path(config.work_dir) / f"{step}"
                if not (config.save_model_checkpoint and config.save_frequency > 0 and step % config.save_frequency == 0):
                    ckpt.save_checkpoint(model, optimizer, step, trained_tokens, str(out_dir),
                                         extra={"data_loader": data_loader.state_dict()})
                    if log_rank:
                        torch.save(lr_scheduler.state_dict(), out_dir / "scheduler.pt")
                rank_print(f"stop requested by signal: checkpoint written at step {step}, exiting", is_print_rank=log_rank)
                break
        while pending:
            flush_one()
        ckpt.wait()                      # an asynchronous save in flight is published before the job reports success
        prof.close()
        watchdog.stop()
    except Exception:
        logger.error("training failed", exc_info=True)
        raise
    finally:
        if run is not None:
            run.finish()
        if monitor is not None:
            try:
                monitor.save_stats()
                if is_logging_rank():
                    monitor.print_summary()
            except Exception:
                pass
        cleanup_distributed()


if __name__ == "__main__":
    main()
