"""Tuning sweep of the CTA-pair GEMM: explicit k-splits and tiny-K runs (epilogue + launch cost) per shape.
Usage: python tools/gemm_sweep.py  ->  gpurun_out/gemm_sweep.jsonl"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaletorch_b200 import ops  # noqa: E402
from scaletorch_b200.ops.gemm import gemm_pair  # noqa: E402
from tools.bench_kernels import timeit  # noqa: E402

os.makedirs("gpurun_out", exist_ok=True)
out = open("gpurun_out/gemm_sweep.jsonl", "a")
# wgrad shapes: dW[M=out, N=in] with K = tokens
shapes = [("0.6B qkv", 4096, 1024, 4096), ("0.6B o", 1024, 2048, 4096), ("0.6B gate_up", 6144, 1024, 4096),
          ("0.6B down", 1024, 3072, 4096), ("8B-tp2 down", 4096, 6144, 4096), ("sq4096", 4096, 4096, 4096)]
for name, M, N, K in shapes:
    for kk in (64, K):
        a = torch.randn(kk, M, device="cuda").bfloat16()
        b = torch.randn(kk, N, device="cuda").bfloat16()
        o = torch.zeros(M, N, device="cuda")
        rec = {"shape": name, "M": M, "N": N, "K": kk}
        for s in (1, 2, 3, 4):
            if kk == 64 and s > 1:
                continue
            rec[f"pair_s{s}_us"] = round(1e3 * timeit(lambda: gemm_pair(a, b, o, a_mn=True, b_mn=True, bn=256, splits=s), iters=20), 2)
        for s in (1, 2):
            rec[f"pair128_s{s}_us"] = round(1e3 * timeit(lambda: gemm_pair(a, b, o, a_mn=True, b_mn=True, bn=128, splits=s), iters=20), 2)
        rec["pair_auto_us"] = round(1e3 * timeit(lambda: gemm_pair(a, b, o, a_mn=True, b_mn=True), iters=20), 2)
        rec["pair_acc_us"] = round(1e3 * timeit(lambda: gemm_pair(a, b, o, a_mn=True, b_mn=True, accumulate=True, bn=256, splits=1), iters=20), 2)
        rec["single256_us"] = round(1e3 * timeit(lambda: ops.gemm(a, b, a_mn=True, b_mn=True, out=o, force_bn=256), iters=20), 2)
        rec["single128_us"] = round(1e3 * timeit(lambda: ops.gemm(a, b, a_mn=True, b_mn=True, out=o, force_bn=128), iters=20), 2)
        rec["cublas_us"] = round(1e3 * timeit(lambda: torch.matmul(a.t(), b), iters=20), 2)
        ob = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        ak = torch.randn(M, kk, device="cuda").bfloat16()
        bk = torch.randn(N, kk, device="cuda").bfloat16()
        rec["pair_fprop_bf16out_us"] = round(1e3 * timeit(lambda: gemm_pair(ak, bk, ob, bn=256, splits=1), iters=20), 2)
        print(json.dumps(rec), flush=True)
        out.write(json.dumps(rec) + "\n")
