#!/usr/bin/env python
"""Quick suite through the *trainer CLI* (``tools/train.py``) instead of ``bench.py``: exercises config parsing, the data
loader, the metrics line and the per-rank ``performance_logs_*.json`` files the way a user run does (reference:
``scripts/benchmark_all.py``, which regex-parses the console line; the same line format is kept by
``scaletorch_b200/trainer/metrics.py``)."""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LINE = re.compile(r"Step:\s*(\d+).*?Loss:\s*([\d.]+).*?Tokens/s:\s*([\d.]+[KMBT]?).*?Tokens/s/GPU:\s*([\d.]+[KMBT]?).*?MFU:\s*([\d.]+)%")

SUITE = [  # (label, model, tp, pp, cp, ep, dp, mbs, ga, seq, extra)
    ("0.6B DP", "Qwen3-0.6B", 1, 1, 1, 1, None, 2, 1, 2048, []),
    ("0.6B TP2", "Qwen3-0.6B", 2, 1, 1, 1, None, 2, 1, 2048, ["--sequence_parallel", "true"]),
    ("0.6B CP2", "Qwen3-0.6B", 1, 1, 2, 1, None, 1, 1, 4096, []),
    ("0.6B PP2", "Qwen3-0.6B", 1, 2, 1, 1, None, 2, 4, 2048, []),
    ("1.7B TP2-CP2", "Qwen3-1.7B", 2, 1, 2, 1, None, 1, 1, 4096, ["--sequence_parallel", "true"]),
    ("30B-A3B EP", "Qwen3-30B-A3B", 1, 1, 1, None, 1, 1, 1, 4096, []),
]


def unit(x: str) -> float:
    mult = {"K": 1e3, "M": 1e6, "B": 1e9, "T": 1e12}
    return float(x[:-1]) * mult[x[-1]] if x[-1] in mult else float(x)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--out", default=os.path.join(ROOT, "results", "benchmark_all.json"))
    a = ap.parse_args()
    results = []
    for label, model, tp, pp, cp, ep, dp, mbs, ga, seq, extra in SUITE:
        ep = ep or (a.gpus // (tp * pp * cp))
        dp = dp or (a.gpus // (tp * pp * cp * ep))
        if tp * pp * cp * ep * dp != a.gpus:
            continue
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={a.gpus}", "--master-addr", "127.0.0.1",
               "--master-port", "29555", os.path.join(ROOT, "tools", "train.py"), "--model_name_or_path", model, "--synthetic_data", "true",
               "--tensor_parallel_size", str(tp), "--pipeline_parallel_size", str(pp), "--context_parallel_size", str(cp),
               "--expert_parallel_size", str(ep), "--data_parallel_size", str(dp), "--micro_batch_size", str(mbs),
               "--gradient_accumulation_steps", str(ga), "--sequence_length", str(seq), "--total_train_steps", str(a.steps)] + extra
        p = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT)
        hits = LINE.findall(p.stdout + p.stderr)
        steady = hits[2:] or hits                         # skip two warm-up steps like the reference
        rec = {"label": label, "rc": p.returncode}
        if steady:
            rec.update(tokens_per_s=sum(unit(h[2]) for h in steady) / len(steady), mfu=sum(float(h[4]) for h in steady) / len(steady),
                       last_loss=float(steady[-1][1]))
        results.append(rec)
        print(rec, flush=True)
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(results, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
