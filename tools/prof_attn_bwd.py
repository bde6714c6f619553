"""Times the attention backward kernels alone (no autograd): python tools/prof_attn_bwd.py [B S Hq Hkv]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaletorch_b200.ops import attention as A  # noqa: E402
from tools.bench_kernels import timeit  # noqa: E402

B, S, Hq, Hkv = (int(x) for x in sys.argv[1:5]) if len(sys.argv) >= 5 else (2, 2048, 16, 8)
D = 128
torch.manual_seed(0)
q = torch.randn(B, S, Hq, D, device="cuda").bfloat16()
k = torch.randn(B, S, Hkv, D, device="cuda").bfloat16()
v = torch.randn(B, S, Hkv, D, device="cuda").bfloat16()
scale = D ** -0.5
out, lse = A._flash_fwd(q, k, v, True, scale)
do = torch.randn_like(out)
dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
dq32 = torch.zeros(q.shape, device="cuda", dtype=torch.float32)
flops = 2.5 * 4.0 * B * Hq * S * S * D / 2
t2 = timeit(lambda: A._flash_bwd(do, q, k, v, out, lse, dq, dk, dv, True, scale), iters=20)
tf = timeit(lambda: A._flash_bwd_fused(do, q, k, v, out, lse, dq32, dk, dv, True, scale), iters=20)
print(f"two-kernel {t2 * 1e3:.1f} us ({flops / t2 / 1e9:.0f} TFLOP/s)   fused(+delta) {tf * 1e3:.1f} us ({flops / tf / 1e9:.0f} TFLOP/s)")
