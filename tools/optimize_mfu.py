#!/usr/bin/env python
"""Sweep (batch size, sequence length, gradient checkpointing) and report MFU with ``6N + 12·L·H·d·S``
(reference: ``tools/optimize_mfu.py:27-103``; its ``torch.compile`` axis is dropped -- no tracing compiler on the hot path)."""
import argparse
import itertools
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from bench_single import benchmark  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="Qwen3-0.6B")
    ap.add_argument("--batch_sizes", default="1,2,4,8")
    ap.add_argument("--seq_lens", default="2048,4096,8192")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--cpu", action="store_true", help="host dry run of the sweep (fp32, wall clock)")
    a = ap.parse_args()
    rows = []
    for bs, sl, gc_ in itertools.product(map(int, a.batch_sizes.split(",")), map(int, a.seq_lens.split(",")), (False, True)):
        try:
            ms, mem, tps = benchmark(a.model, bs, sl, a.steps, gradient_checkpointing=gc_, cpu=a.cpu)
            rows.append((bs, sl, gc_, ms, tps, mem))
        except torch.cuda.OutOfMemoryError:
            rows.append((bs, sl, gc_, None, None, None))
            torch.cuda.empty_cache()
    print("\nBS   SEQ    GC     step_ms     tok/s     peak_GB")
    for bs, sl, gc_, ms, tps, mem in rows:
        print(f"{bs:<4d} {sl:<6d} {str(gc_):<6s} " + ("OOM" if ms is None else f"{ms:9.1f} {tps:10.0f} {mem:10.2f}"))


if __name__ == "__main__":
    main()
