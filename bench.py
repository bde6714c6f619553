#!/usr/bin/env python
"""Flagship benchmark: Qwen3-0.6B, DP over N GPUs, bf16, seq 2048, micro-batch 2 per GPU.

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the unmodified reference from baseline/_ref

Prints ONE JSON line on rank 0 (see README "Benchmark contract").  ``value`` is whole-job tokens/s for full
optimiser steps (forward + backward + gradient all-reduce + clip + AdamW), device-timed with CUDA events
between barriers, max over ranks.  ``e2e`` is the same metric through the public training API including,
every step, the pinned-host → device copy of that step's token batch and a device → host read of the loss.
Synthetic tokens, random-init weights (no network on the box).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
BASELINE_TOK_S_PER_DEVICE = {  # BASELINE.md §2 (Ascend 910B, README.md:48) -- context, other hardware
    "Qwen3-0.6B": 16422.0,
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--model", default="Qwen3-0.6B")
    ap.add_argument("--micro_batch_size", type=int, default=2)
    ap.add_argument("--seq_len", type=int, default=2048)
    ap.add_argument("--grad_accum", type=int, default=1)
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--cp", type=int, default=1)
    ap.add_argument("--pp", type=int, default=1)
    ap.add_argument("--ep", type=int, default=1)
    ap.add_argument("--sequence_parallel", action="store_true")
    ap.add_argument("--gradient_checkpointing", action="store_true")
    ap.add_argument("--skip_e2e", action="store_true")
    ap.add_argument("--cp_zigzag", type=int, default=1, help="load-balanced causal context parallelism (ours)")
    ap.add_argument("--cuda_graph", type=int, default=1, help="capture the whole step in a CUDA graph when supported (ours)")
    return ap.parse_args()


class ClockSampler:
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 7:
                    self.rows.append(parts)
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._t is not None:
            self._t.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "", 1).isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "", 1).isdigit()]
        reasons = set()
        for r in self.rows:
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(self.rows)}


def fused_mc(gw) -> int:
    """NVLS multicast address of the gradient window even when the default transport choice disabled it."""
    from scaletorch_b200.runtime.fused_ops import _multicast_ptr

    return _multicast_ptr(gw.h)


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def timed_region(torch, dist, world, fn, steps):
    """barrier+sync, K × fn() between CUDA events, barrier+sync; returns ms (max over ranks)."""
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def build_config_dict(args, world):
    dp = world // (args.tp * args.cp * args.pp * args.ep)
    par = "".join(f"{k}{v}" for k, v in (("tp", args.tp), ("cp", args.cp), ("pp", args.pp), ("ep", args.ep)) if v > 1)
    par = (par + f"dp{dp}") if par else f"dp{dp}"
    replicas = dp * args.ep
    return dp, {
        "model": args.model, "global_batch": replicas * args.micro_batch_size * args.grad_accum, "seq_len": args.seq_len,
        "micro_batch_size": args.micro_batch_size, "grad_accum": args.grad_accum, "parallelism": par,
        "sequence_parallel": bool(args.sequence_parallel), "gradient_checkpointing": bool(args.gradient_checkpointing),
        "optimizer": "adamw + global-norm clip 1.0",
        "l2_policy": "inputs larger than L2: every step streams >1 GB of weights/activations (126 MB L2); no explicit flush",
    }


# ----------------------------------------------------------------------------------------- ours
def run_ours(args):
    import torch
    import torch.distributed as dist

    sys.path.insert(0, ROOT)
    from scaletorch_b200 import _native
    from scaletorch_b200.data import MicroBatchDataLoader
    from scaletorch_b200.trainer import (ScaleTorchArguments, create_model, create_optimizer,
                                         initialize_distributed_training, training_step)
    from scaletorch_b200.utils.misc import get_mfu, get_num_params

    rank, local_rank, world = dist_env()
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run for N>1")
    dp, cfgd = build_config_dict(args, world)
    cfg = ScaleTorchArguments(
        model_name_or_path=args.model, micro_batch_size=args.micro_batch_size, sequence_length=args.seq_len,
        gradient_accumulation_steps=args.grad_accum, tensor_parallel_size=args.tp, context_parallel_size=args.cp,
        pipeline_parallel_size=args.pp, expert_parallel_size=args.ep, data_parallel_size=dp,
        sequence_parallel=args.sequence_parallel, gradient_checkpointing=args.gradient_checkpointing,
        cp_zigzag=bool(args.cp_zigzag), synthetic_data=True, dtype="bfloat16", learning_rate=3e-4, max_grad_norm=1.0, seed=1)
    _, _, _, device = initialize_distributed_training(cfg)
    _native.load(required=True)
    model, model_config = create_model(cfg, torch.bfloat16, device)
    model.train()
    optimizer = create_optimizer(model, cfg, device)
    num_params = get_num_params(model)
    loader = MicroBatchDataLoader(args.micro_batch_size, args.seq_len, gradient_accumulation_steps=args.grad_accum,
                                  synthetic=True, vocab_size=model_config.vocab_size, seed=1 + rank, pin_memory=True,
                                  num_samples=max(4096, (args.steps + args.warmup + 2) * 4 * args.micro_batch_size * world))

    class DeviceBatches:
        """Pre-staged device-resident batches for the device-timed region (a fresh batch every step)."""

        def __init__(self, n):
            self.b = []
            for _ in range(n):
                x = next(loader)
                self.b.append({k: (v.to(device) if torch.is_tensor(v) else v) for k, v in x.items()})
            self.i = 0
            self.gradient_accumulation_steps = args.grad_accum

        def __next__(self):
            x = self.b[self.i % len(self.b)]
            self.i += 1
            return x

    from scaletorch_b200.trainer import GraphedTrainStep

    graphed = None
    if args.cuda_graph and GraphedTrainStep.supported(cfg, optimizer, model):
        graphed = GraphedTrainStep(model, optimizer, cfg, device, warmup=2)
    launches_per_graph = [0]
    per_kernel_graph: dict = {}

    def step_with(src):
        if graphed is not None:
            was_captured = graphed.graph is not None
            c0, k0 = _native.launch_count, dict(_native.launch_counts)
            loss = graphed(src)
            if not was_captured and graphed.graph is not None:
                launches_per_graph[0] = _native.launch_count - c0      # kernels recorded into the graph
                per_kernel_graph.update({k: v - k0.get(k, 0) for k, v in _native.launch_counts.items() if v > k0.get(k, 0)})
            return loss
        optimizer.zero_grad()
        if hasattr(model, "reset"):
            model.reset()
        loss = training_step(model, src, device, torch.bfloat16, cfg)
        optimizer.step()
        return loss

    staged = DeviceBatches((args.steps + args.warmup) * args.grad_accum)
    for _ in range(args.warmup):
        step_with(staged)
    while graphed is not None and graphed.graph is None and not graphed.failed:   # the capture never falls into the timed region
        step_with(staged)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    if os.environ.get("SCALETORCH_B200_PROFILE"):
        # diagnosis only (kernel-time table of rank 0 -> file); numbers of a profiled run are never reported
        from torch.profiler import ProfilerActivity, profile

        with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
            for _ in range(2):
                step_with(staged)
            torch.cuda.synchronize()
        if rank == 0:
            with open(os.environ["SCALETORCH_B200_PROFILE"], "w") as f:
                f.write(prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=90))
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    if os.environ.get("SCALETORCH_B200_DP_SWEEP"):
        # diagnosis: time the same model under several (transport, CTA budget) settings of the gradient all-reduce
        dp = model if hasattr(model, "bucket_manager") else None
        gw = getattr(getattr(dp, "store", None), "grad_window", None)
        if dp is not None and gw is not None:
            mc = gw.mc_ptr or fused_mc(gw)
            for spec in os.environ["SCALETORCH_B200_DP_SWEEP"].split(","):
                kind, budget = spec.split(":")
                gw.mc_ptr = mc if kind == "nvls" else 0
                for b in dp.bucket_manager.buckets:
                    b.sm_budget = int(budget)
                if graphed is not None:            # the transport and the CTA budget are baked into the graph: re-capture
                    graphed.graph = None
                for _ in range(2):
                    step_with(staged)
                t = timed_region(torch, dist, world, lambda: step_with(staged), args.steps) / args.steps
                if rank == 0:
                    print(f"DP_SWEEP {kind} budget {budget}: {t:.3f} ms/step", flush=True)
            gw.mc_ptr = mc
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    l0, k0 = _native.launch_count, dict(_native.launch_counts)
    ms = timed_region(torch, dist, world, lambda: step_with(staged), args.steps)
    launches = _native.launch_count - l0
    per_kernel = {k: v - k0.get(k, 0) for k, v in _native.launch_counts.items() if v > k0.get(k, 0)}
    if graphed is not None and graphed.graph is not None:
        launches = launches_per_graph[0] * args.steps              # replayed from the graph: same kernels every step
        per_kernel = {k: v * args.steps for k, v in per_kernel_graph.items()}
    comm_kernels = ("stb_allreduce_f32", "stb_reduce_scatter_f32", "stb_adamw_shard_bcast", "stb_param_bcast", "stb_allreduce_tiny_f32",
                    "stb_embedding_bwd_peers", "stb_ag_gemm_bf16", "stb_gemm_rs_bf16", "stb_allreduce_bf16",
                    "stb_allreduce_oneshot_bf16", "stb_ep_push", "stb_ep_pull", "stb_ep_exchange_counts", "stb_symm_barrier")
    # which transport the gradient synchronisation used: NVLS multimem / peer loads-stores / c10d-NCCL
    store = getattr(model, "store", None) or getattr(model, "_stb_store", None)
    gw = getattr(store, "grad_window", None)
    dp_transport = None
    if world > 1 and dp * args.cp * (args.ep if args.ep > 1 else 1) > 1:
        dp_transport = "nccl" if gw is None else ("nvls" if gw.mc_ptr else "p2p")
    clocks = sampler.stop() if rank == 0 else {}

    # end-to-end: public API, per-step pinned H2D of the batch + D2H read of the loss
    e2e = None
    if not args.skip_e2e:
        from scaletorch_b200.trainer import AsyncScalarReader

        losses = []
        reader = AsyncScalarReader(slots=4)

        def e2e_step():
            # every step: pinned H2D of the batch (inside the loader / train step), D2H copy of the loss into pinned
            # host memory.  The host consumes step i's loss after enqueueing step i+1 (one-step lag), as the trainer
            # CLI does, so the device queue never drains; all K copies land inside the timed region.
            reader.push(step_with(loader))
            if len(reader) > 1:
                losses.append(reader.pop())

        e2e_step()
        ms_e2e = timed_region(torch, dist, world, e2e_step, args.steps)
        losses.extend(reader.drain())
        tokens = cfgd["global_batch"] * args.seq_len
        per_rank_tokens = args.micro_batch_size * args.grad_accum * (args.seq_len // args.cp)
        e2e = {"value": tokens * args.steps / (ms_e2e / 1e3), "unit": "tokens/s", "ms_per_step": ms_e2e / args.steps,
               "h2d_bytes_per_step": per_rank_tokens * 8 * 3, "d2h_bytes_per_step": 4, "last_loss": losses[-1]}

    # expert parallelism: (token, expert) pairs dropped because a receive region was full, summed over layers and ranks
    # (0 by construction unless a capacity factor below the EP size was asked for)
    ep_dropped = None
    if args.ep > 1:
        tot = torch.zeros(1, dtype=torch.float64, device=device)
        for m in model.modules():
            d = getattr(m, "dropped_pairs", None)
            if d is not None:
                tot += d.double().sum()
        dist.all_reduce(tot)
        ep_dropped = int(tot.item())
    if rank == 0:
        tokens = cfgd["global_batch"] * args.seq_len
        value = tokens * args.steps / (ms / 1e3)
        base = BASELINE_TOK_S_PER_DEVICE.get(args.model)
        out = {
            "metric": "tokens/s (whole job; tokens/s/GPU = value / n_gpus) + MFU", "value": value, "unit": "tokens/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": (value / (base * world)) if base else None,
            "dtype": "bf16", "data": "synthetic tokens, random-init weights", "impl": "ours", "config": cfgd,
            "tokens_per_s_per_gpu": value / world,
            "mfu_pct_of_nominal_2.25PF": get_mfu(value / world, num_params, model_config, 2.25e15, args.seq_len),
            "num_params": num_params, "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
            "comm_kernel_launches": {k: per_kernel.get(k, 0) for k in comm_kernels if per_kernel.get(k, 0)},
            "dp_transport": dp_transport, "optimizer": type(optimizer).__name__, "ep_dropped_pairs": ep_dropped,
            "ep_capacity_factor": os.environ.get("SCALETORCH_B200_EP_CAPACITY", "dropless") if args.ep > 1 else None,
            "cuda_graph": bool(graphed is not None and graphed.graph is not None),
            "native_lib": str(_native.lib_path()),
        }
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        if peaks.get("bf16_tflops_sustained"):
            out["mfu_pct_of_measured_sustained"] = get_mfu(value / world, num_params, model_config,
                                                           peaks["bf16_tflops_sustained"] * 1e12, args.seq_len)
        print(json.dumps(out), flush=True)
    if graphed is not None:
        graphed.release()                           # graphs with collective nodes go before the communicator does
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------ reference
def run_reference(args):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    sys.path.insert(0, os.path.join(ROOT, "baseline"))  # run_reference.py, install_reference.py + the colorama stand-in
    rank, local_rank, world = dist_env()
    if not os.path.isdir(os.path.join(ref_dir, "scaletorch")):
        # lazily install wherever /root/reference is visible (local rank 0 installs, the others wait for the files)
        import install_reference

        if local_rank == 0:
            install_reference.ensure()
        else:
            for _ in range(600):
                if install_reference.installed():
                    break
                time.sleep(0.5)
    if not os.path.isdir(os.path.join(ref_dir, "scaletorch")):
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref/scaletorch not installed and /root/reference not visible"}))
        return
    try:
        import run_reference

        run_reference.main(args, ROOT, ref_dir, ClockSampler, timed_region, build_config_dict, dist_env,
                           BASELINE_TOK_S_PER_DEVICE)
    except Exception as e:  # the reference has known defects at this snapshot
        import traceback

        traceback.print_exc()
        rank = int(os.environ.get("RANK", 0))
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": f"{type(e).__name__}: {str(e)[:300]}"}))


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
