#!/usr/bin/env python
"""Will this model fit?  Per-GPU memory plan of a (model, layout, batch) choice against the 180 GB of a B200.

    python tools/plan_memory.py --model Qwen3-32B --tp 4 --pp 2 --micro_batch_size 1 --sequence_length 4096 --ga 8 --sp
    python tools/plan_memory.py --model Qwen3-30B-A3B --ep 8 --sequence_length 4096 --json

No GPU and no process group are needed: the model is built on the meta device and its parameters are cut with the SAME
functions the trainer uses (``utils/sharding.py::shard_tensor`` for TP / EP, ``distribute_layers`` for PP), so the static
part -- parameters (bf16), the flat fp32 gradient buffer, AdamW state, symmetric-memory windows -- is exact arithmetic on
real shapes.  The activation part is an ESTIMATE: the tensors each decoder layer's autograd nodes keep for backward are
enumerated from the layer's structure (residual stream, normed inputs, q|k|v before and after QK-norm + RoPE, attention
output and LSE, gate|up output, SwiGLU output; MoE: permuted tokens and expert intermediates), times the micro-batches in
flight on the stage.  It has not been calibrated against a measured peak; treat it as a first-order figure and leave
head-room for the allocator.

The reference has no such tool; its README tables list memory per configuration after the fact.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from scaletorch_b200.parallel.context import single_device_context  # noqa: E402
from scaletorch_b200.parallel.pipeline_parallel.pipeline_parallel import distribute_layers  # noqa: E402
from scaletorch_b200.trainer.config import resolve_model_config  # noqa: E402
from scaletorch_b200.trainer.model_builder import model_class_for  # noqa: E402
from scaletorch_b200.utils.sharding import shard_tensor  # noqa: E402

GB = 1e9


def stage_parameters(cfg, tp: int, pp: int, ep: int, stage: int) -> dict:
    """Parameter counts held by ONE rank of pipeline stage ``stage`` (tp_rank = ep_rank = 0; every rank of a stage holds
    the same amount): ``{"dense": n, "expert": n}``."""
    with torch.device("meta"):
        model = model_class_for(cfg)(cfg, single_device_context())
    layers = set(distribute_layers(cfg.num_hidden_layers, pp, stage))
    tied = bool(getattr(cfg, "tie_word_embeddings", False))
    dense = expert = 0
    for name, p in model.named_parameters():
        if name.startswith("decoder_layers."):
            if int(name.split(".")[1]) not in layers:
                continue
        elif name.startswith("embedding."):
            # the first stage looks tokens up; with tied weights the last stage holds its own copy for the LM head
            if not (stage == 0 or (tied and stage == pp - 1)):
                continue
        elif name.startswith(("final_norm.", "final_proj.")):
            if stage != pp - 1:
                continue
        n = shard_tensor(name, p, cfg, tp, 0, ep, 0).numel()
        if ".experts." in name:
            expert += n
        else:
            dense += n
    return {"dense": dense, "expert": expert}


def ep_rows_cap(cfg, tokens: int, ep: int, capacity_factor: float | None) -> int:
    """Rows of the expert-major padded layout on one EP rank (``fused_ep.fused_moe_ep``): dropless = the worst case
    ``ep x T x k`` unless a capacity factor bounds it; every expert's segment is padded to 128 rows."""
    k = cfg.num_experts_per_tok
    cf = float(ep) if capacity_factor is None else min(float(capacity_factor), float(ep))
    e_local = cfg.num_experts // ep
    return (int(cf * tokens * k) + e_local * 127 + 127) // 128 * 128


def layer_activation_bytes(cfg, tokens: int, tp: int, sp: bool, is_moe: bool, ep: int = 1,
                           ep_capacity_factor: float | None = None) -> int:
    """Bytes one decoder layer keeps for backward for ``tokens`` local tokens (bf16 unless noted)."""
    H = cfg.hidden_size
    hd = getattr(cfg, "head_dim", None) or H // cfg.num_attention_heads
    nq, nkv = cfg.num_attention_heads // tp, max(1, cfg.num_key_value_heads // tp)
    t_res = tokens // tp if sp else tokens                    # residual stream and normed inputs are sequence-sharded under SP
    b = 0
    b += 2 * t_res * H * 2                                    # residual after each of the two adds
    b += 2 * t_res * H * 2                                    # the two normed GEMM inputs
    b += 2 * t_res * 4                                        # rstd of the two norms (fp32)
    b += 2 * tokens * (nq + 2 * nkv) * hd * 2                 # q|k|v before and after QK-norm + RoPE
    b += tokens * nq * hd * 2 + tokens * nq * 4               # attention output, LSE (fp32)
    if is_moe:
        k, inter = cfg.num_experts_per_tok, cfg.moe_intermediate_size
        # with expert parallelism the row-padded tensors have the STATIC capacity of the receive region, not T x k
        rows = ep_rows_cap(cfg, tokens, ep, ep_capacity_factor) if ep > 1 else tokens * k
        b += tokens * cfg.num_experts * 4                     # router probabilities (fp32)
        b += rows * H * 2 * 2                                 # permuted expert inputs and outputs
        b += rows * 3 * inter * 2                             # gate|up output + SwiGLU output
    else:
        inter = cfg.intermediate_size // tp
        b += tokens * 3 * inter * 2                           # gate|up output + SwiGLU output
    return b


def plan(model: str, tp=1, pp=1, cp=1, dp=1, ep=1, micro_batch_size=1, sequence_length=4096, ga=1, sp=False,
         gradient_checkpointing=False, zero_stage=None, window_mb=1024, hbm_gb=180.0, ep_capacity_factor=None) -> dict:
    cfg = resolve_model_config(model)
    if cfg.num_attention_heads % tp or cfg.num_key_value_heads % tp:
        raise SystemExit(f"tp={tp} must divide the head counts ({cfg.num_attention_heads} / {cfg.num_key_value_heads})")
    if sequence_length % cp:
        raise SystemExit(f"cp={cp} must divide the sequence length")
    E = getattr(cfg, "num_experts", 0) or 0
    if ep > 1 and (E == 0 or E % ep):
        raise SystemExit(f"ep={ep} needs a MoE model whose expert count ({E}) it divides")
    tokens = micro_batch_size * sequence_length // cp
    dense_sync = dp * cp * ep                                  # ranks that average dense gradients
    expert_sync = dp * cp
    peer_zero = pp == 1 and dense_sync > 1                      # PeerShardedAdamW: the default when there is no pipeline split
    step = getattr(cfg, "decoder_sparse_step", 1) or 1
    mlp_only = getattr(cfg, "mlp_only_layers", None) or []
    stages = []
    for s in range(pp):
        n = stage_parameters(cfg, tp, pp, ep, s)
        P = n["dense"] + n["expert"]
        if zero_stage == 1:
            opt = 8 * n["dense"] / dense_sync + 8 * n["expert"] / expert_sync
        elif peer_zero and zero_stage is None:
            opt = 8 * n["dense"] / dense_sync + 8 * n["expert"]
        else:
            opt = 8 * P
        layer_ids = distribute_layers(cfg.num_hidden_layers, pp, s)
        per_layer = [layer_activation_bytes(cfg, tokens, tp, sp, bool(E) and i not in mlp_only and (i + 1) % step == 0, ep,
                                            ep_capacity_factor) for i in layer_ids]
        in_flight = min(pp - s, ga) if pp > 1 else 1           # 1F1B: stage s holds up to pp - s micro-batches
        t_res = tokens // tp if sp else tokens
        if gradient_checkpointing:
            act = in_flight * len(layer_ids) * (2 * t_res * cfg.hidden_size * 2) + (max(per_layer) if per_layer else 0)
        else:
            act = in_flight * sum(per_layer)
        logits = tokens * (cfg.vocab_size // tp) * 2 if s == pp - 1 else 0
        windows = 0
        if tp > 1:
            windows += window_mb << 20                          # TP group window (all-gather / reduce-scatter / all-reduce regions)
        if cp > 1:
            windows += window_mb << 20                          # CP group window (K/V blocks, fp32 dK/dV accumulators)
        if ep > 1:
            windows += 2 * ep_rows_cap(cfg, tokens, ep, ep_capacity_factor) * cfg.hidden_size * 2    # receive + return regions
        total = 2 * P + 4 * P + opt + act + logits + windows
        stages.append({"stage": s, "layers": len(layer_ids), "params": P, "params_expert": n["expert"],
                       "bytes": {"params_bf16": 2 * P, "grads_fp32": 4 * P, "adamw_state": int(opt), "activations_est": int(act),
                                 "logits": int(logits), "symmetric_windows": int(windows), "total": int(total)},
                       "micro_batches_in_flight": in_flight, "fits": total <= hbm_gb * GB,
                       "fraction_of_hbm": round(total / (hbm_gb * GB), 3)})
    worst = max(stages, key=lambda st: st["bytes"]["total"])
    return {"model": model, "layout": {"tp": tp, "pp": pp, "cp": cp, "dp": dp, "ep": ep, "world": tp * pp * cp * dp * ep},
            "micro_batch_size": micro_batch_size, "sequence_length": sequence_length, "gradient_accumulation": ga,
            "sequence_parallel": sp, "gradient_checkpointing": gradient_checkpointing,
            "optimizer": ("zero1 (c10d)" if zero_stage == 1 else "zero1 on peer memory (dense sharded)" if peer_zero and zero_stage is None
                          else "replicated AdamW"),
            "ep_capacity_factor": (None if ep == 1 else ("dropless" if ep_capacity_factor is None else ep_capacity_factor)),
            "hbm_gb": hbm_gb, "stages": stages, "worst_stage": worst["stage"], "fits": all(st["fits"] for st in stages)}


def suggest(model: str, gpus: int, **kw) -> list[dict]:
    """Every layout of ``gpus`` GPUs (powers of two per axis) that is legal for the model and fits, most data-parallel
    first (data parallelism is the cheapest axis on an NVSwitch node: docs/parallelism.md), then by memory.  A list of
    candidates to measure, not a performance prediction."""
    cfg = resolve_model_config(model)
    E = getattr(cfg, "num_experts", 0) or 0
    sizes = [n for n in (1, 2, 4, 8, 16, 32, 64) if n <= gpus]
    found = []
    for tp in sizes:
        if cfg.num_attention_heads % tp or cfg.num_key_value_heads % tp:
            continue
        for pp in sizes:
            if pp > cfg.num_hidden_layers:
                continue
            for cp in sizes:
                if kw.get("sequence_length", 4096) % (2 * cp if cp > 1 else 1):
                    continue
                for ep in (sizes if E else [1]):
                    if E and E % ep:
                        continue
                    used = tp * pp * cp * ep
                    if used > gpus or gpus % used:
                        continue
                    dp = gpus // used
                    try:
                        p = plan(model, tp=tp, pp=pp, cp=cp, dp=dp, ep=ep, sp=tp > 1, **kw)
                    except SystemExit:
                        continue
                    if p["fits"]:
                        worst = max(st["fraction_of_hbm"] for st in p["stages"])
                        # combinations whose training equality is pinned by tests/test_parallel_cpu.py (gloo) today
                        axes = {n for n, v in (("tp", tp), ("pp", pp), ("cp", cp), ("ep", ep), ("dp", dp)) if v > 1}
                        # pairs, triples, TP x CP x DP x EP and all five axes together are pinned by tests
                        covered = len(axes) <= 3 or axes == {"tp", "cp", "dp", "ep"} or len(axes) == 5
                        found.append({"tp": tp, "pp": pp, "cp": cp, "ep": ep, "dp": dp, "worst_fraction_of_hbm": worst,
                                      "optimizer": p["optimizer"], "combination_covered_by_tests": covered})
    found.sort(key=lambda f: (not f["combination_covered_by_tests"], -f["dp"], f["tp"] * f["pp"] * f["cp"], f["worst_fraction_of_hbm"]))
    return found


def render(p: dict) -> str:
    lay = p["layout"]
    out = [f"{p['model']}  tp{lay['tp']} pp{lay['pp']} cp{lay['cp']} dp{lay['dp']} ep{lay['ep']} ({lay['world']} GPUs)  "
           f"mbs {p['micro_batch_size']} x S {p['sequence_length']} x GA {p['gradient_accumulation']}  "
           f"SP={p['sequence_parallel']} GC={p['gradient_checkpointing']}  optimiser: {p['optimizer']}",
           f"{'stage':>5} {'layers':>6} {'params':>9} {'bf16':>8} {'grads':>8} {'adamw':>8} {'act(est)':>9} {'logits':>8} "
           f"{'windows':>8} {'total':>9} {'of HBM':>7}"]
    for st in p["stages"]:
        b = st["bytes"]
        out.append(f"{st['stage']:>5d} {st['layers']:>6d} {st['params'] / 1e9:>8.2f}B {b['params_bf16'] / GB:>7.1f}G {b['grads_fp32'] / GB:>7.1f}G "
                   f"{b['adamw_state'] / GB:>7.1f}G {b['activations_est'] / GB:>8.1f}G {b['logits'] / GB:>7.1f}G "
                   f"{b['symmetric_windows'] / GB:>7.1f}G {b['total'] / GB:>8.1f}G {st['fraction_of_hbm'] * 100:>6.1f}%")
    out.append(("FITS" if p["fits"] else "DOES NOT FIT") + f" in {p['hbm_gb']:.0f} GB per GPU (activations are an estimate; "
               "leave head-room for the allocator and NCCL buffers)")
    return "\n".join(out)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--model", required=True, help="model name or directory, as given to --model_name_or_path")
    for axis in ("tp", "pp", "cp", "dp", "ep"):
        ap.add_argument(f"--{axis}", type=int, default=1)
    ap.add_argument("--micro_batch_size", type=int, default=1)
    ap.add_argument("--sequence_length", type=int, default=4096)
    ap.add_argument("--ga", type=int, default=1, help="gradient accumulation steps (micro-batches per step; bounds the 1F1B in-flight count)")
    ap.add_argument("--sp", action="store_true", help="sequence parallelism (with tp > 1)")
    ap.add_argument("--gradient_checkpointing", action="store_true")
    ap.add_argument("--zero_stage", type=int, default=None, choices=(0, 1), help="default: what the trainer picks for the layout")
    ap.add_argument("--ep_capacity_factor", type=float, default=None,
                    help="receive capacity of the EP path as a multiple of tokens x top_k (default: dropless = ep)")
    ap.add_argument("--window_mb", type=int, default=int(os.environ.get("SCALETORCH_B200_SYMM_WINDOW_MB", "1024")))
    ap.add_argument("--hbm_gb", type=float, default=180.0)
    ap.add_argument("--json", action="store_true")
    ap.add_argument("--suggest", action="store_true", help="list every layout of --gpus GPUs that fits (ignores the axis flags)")
    ap.add_argument("--gpus", type=int, default=8)
    a = ap.parse_args(argv)
    try:
        resolve_model_config(a.model)
    except RuntimeError as e:
        vendored = sorted(os.listdir(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "configs")))
        raise SystemExit(f"plan_memory.py: {str(e).splitlines()[0]}\nvendored configs: {', '.join(vendored)}")
    if a.suggest:
        found = suggest(a.model, a.gpus, micro_batch_size=a.micro_batch_size, sequence_length=a.sequence_length, ga=a.ga,
                        gradient_checkpointing=a.gradient_checkpointing, zero_stage=a.zero_stage, window_mb=a.window_mb,
                        hbm_gb=a.hbm_gb, ep_capacity_factor=a.ep_capacity_factor)
        if a.json:
            print(json.dumps(found))
        else:
            print(f"{a.model} on {a.gpus} GPUs, mbs {a.micro_batch_size} x S {a.sequence_length}: {len(found)} layouts fit in {a.hbm_gb:.0f} GB")
            for f in found[:24]:
                print(f"  dp{f['dp']:<3d} tp{f['tp']:<2d} pp{f['pp']:<2d} cp{f['cp']:<2d} ep{f['ep']:<2d}  worst stage {f['worst_fraction_of_hbm'] * 100:5.1f}% of HBM   {f['optimizer']}"
                      + ("" if f["combination_covered_by_tests"] else "   (axis combination not covered by the equality tests)"))
        return 0 if found else 1
    p = plan(a.model, a.tp, a.pp, a.cp, a.dp, a.ep, a.micro_batch_size, a.sequence_length, a.ga, a.sp and a.tp > 1,
             a.gradient_checkpointing, a.zero_stage, a.window_mb, a.hbm_gb, a.ep_capacity_factor)
    print(json.dumps(p) if a.json else render(p))
    return 0 if p["fits"] else 1


if __name__ == "__main__":
    sys.exit(main())
