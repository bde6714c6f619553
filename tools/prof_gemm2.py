"""ncu driver: one plain and one CTA-pair GEMM on the same shape.  Usage: python tools/prof_gemm2.py [M N K] [mode]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaletorch_b200 import ops  # noqa: E402
from scaletorch_b200.ops.gemm import gemm_pair  # noqa: E402

M, N, K = (int(x) for x in sys.argv[1:4]) if len(sys.argv) >= 4 else (4096, 4096, 4096)
mode = sys.argv[4] if len(sys.argv) > 4 else "fprop"
a_mn, b_mn = {"fprop": (False, False), "dgrad": (False, True), "wgrad": (True, True)}[mode]
torch.manual_seed(0)
a = torch.randn((K, M) if a_mn else (M, K), device="cuda").bfloat16()
b = torch.randn((K, N) if b_mn else (N, K), device="cuda").bfloat16()
out = torch.empty(M, N, device="cuda", dtype=torch.float32 if mode == "wgrad" else torch.bfloat16)
for _ in range(3):
    ops.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out)
    gemm_pair(a, b, out, a_mn=a_mn, b_mn=b_mn, splits=0)
    ops.gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, force_bn=256)
torch.cuda.synchronize()
print("ok")
