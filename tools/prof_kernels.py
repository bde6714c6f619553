"""Tiny driver for ncu captures: runs each hot kernel a few times on fixed shapes.
Usage: python tools/prof_kernels.py [attn|gemm|all]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaletorch_b200 import ops  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "all"
torch.manual_seed(0)
if what in ("attn", "all"):
    B, S, Hq, Hkv, D = 2, 2048, 16, 8, 128
    q = torch.randn(B, S, Hq, D, device="cuda").bfloat16().requires_grad_()
    k = torch.randn(B, S, Hkv, D, device="cuda").bfloat16().requires_grad_()
    v = torch.randn(B, S, Hkv, D, device="cuda").bfloat16().requires_grad_()
    for _ in range(3):
        o, _ = ops.flash_attention(q, k, v, True)
        o.backward(torch.randn_like(o))
if what in ("gemm", "all"):
    for (M, N, K) in [(4096, 4096, 1024), (4096, 4096, 4096)]:
        a = torch.randn(M, K, device="cuda").bfloat16()
        b = torch.randn(N, K, device="cuda").bfloat16()
        for _ in range(3):
            ops.gemm(a, b)
torch.cuda.synchronize()
print("ok")
