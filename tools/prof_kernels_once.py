"""Launches every single-GPU kernel family ONCE at a model shape -- the program `ncu --set full` is pointed at.

    python tools/prof_kernels_once.py            # plain run first (must exit 0)
    ncu --set full --clock-control none --import-source on -k regex:'stb|gemm|attn|norm|swiglu|ce_|adamw|router' \
        -o gpurun_out/prof_kernels_r2 python tools/prof_kernels_once.py
    python tools/ncu_summary.py gpurun_out/prof_kernels_r2.ncu-rep > profiles/ncu_r2_kernels.md

Shapes: the dense GEMM is the 8B gate‖up projection the round-1 review names (4096 x 12288 x 4096); attention is the
headline Qwen3-0.6B block (B2 S2048 Hq16 Hkv8 D128); the grouped GEMM is one Qwen3-30B-A3B EP8 rank (16 local experts,
H 2048, I 768, 8192 token-expert pairs); the elementwise kernels use the 0.6B shapes.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaletorch_b200 import _native as N  # noqa: E402
from scaletorch_b200 import ops  # noqa: E402

dev = "cuda"
torch.manual_seed(0)


def bf(*shape):
    return torch.randn(*shape, device=dev).bfloat16()


# ---- dense GEMM: fprop / dgrad / wgrad ---------------------------------------------------------------------------
T, Nn, K = 4096, 12288, 4096
x, w = bf(T, K), bf(Nn, K)
y = ops.gemm(x, w)                                            # fprop   [T,K] . [N,K]^T
dy = bf(T, Nn)
dx = ops.gemm(dy, w, b_mn=True)                               # dgrad   [T,N] . [N,K]
dw = torch.empty(Nn, K, device=dev, dtype=torch.float32)
ops.gemm(dy, x, a_mn=True, b_mn=True, out=dw)                 # wgrad   [T,N]^T . [T,K] -> fp32
del x, w, y, dy, dx, dw

# ---- attention forward + backward --------------------------------------------------------------------------------
B, S, Hq, Hkv, D = 2, 2048, 16, 8, 128
q, k, v = (bf(B, S, h, D).requires_grad_() for h in (Hq, Hkv, Hkv))
out, lse = ops.flash_attention(q, k, v, causal=True)
torch.autograd.grad(out, (q, k, v), torch.randn_like(out))
del q, k, v, out, lse

# ---- grouped GEMM (MoE experts) -------------------------------------------------------------------------------------
from scaletorch_b200.ops.grouped_native import grouped_mlp_native  # noqa: E402

E, H, I, R = 16, 2048, 768, 8192
tpe = torch.full((E,), R // E, dtype=torch.int64, device=dev)
tpe[0] += 37
tpe[1] -= 37                                                  # ragged segments
xs = bf(R, H).requires_grad_()
gu_w = (bf(E, 2 * I, H) * 0.02).requires_grad_()
dn_w = (bf(E, H, I) * 0.02).requires_grad_()
ym = grouped_mlp_native(xs, tpe, gu_w, dn_w)
torch.autograd.grad(ym, (xs, gu_w, dn_w), torch.randn_like(ym))
del xs, gu_w, dn_w, ym

# ---- MoE router --------------------------------------------------------------------------------------------------
from scaletorch_b200.ops.moe_routing import route  # noqa: E402

logits = bf(8192, 128).requires_grad_()
wts, idx, pos, counts, aux = route(logits, 8, 0.001, True)
torch.autograd.grad(wts.sum() + aux, logits)

# ---- normalisation / activation / rope ---------------------------------------------------------------------------
T2, H2 = 4096, 1024
x2, r2 = bf(T2, H2).requires_grad_(), bf(T2, H2).requires_grad_()
w2 = torch.ones(H2, device=dev).bfloat16().requires_grad_()
hh, yy = ops.add_rmsnorm(x2, r2, w2, 1e-6)
torch.autograd.grad((yy, hh), (x2, r2, w2), (torch.randn_like(yy), torch.randn_like(hh)))
gu = bf(T2, 2 * 3072).requires_grad_()
act = ops.swiglu(gu)
torch.autograd.grad(act, gu, torch.randn_like(act))
qkv = bf(B * S, (Hq + 2 * Hkv) * D).requires_grad_()
wq, wk = (torch.ones(D, device=dev).bfloat16().requires_grad_() for _ in range(2))
cos, sin = ops.get_cos_sin(4096, D, 1e6, device=dev)
o = ops.qkv_attention(qkv, wq, wk, cos, sin, B, S, Hq, Hkv, D, eps=1e-6)      # qk-norm + rope + attention (+ fused backward)
torch.autograd.grad(o, (qkv, wq, wk), torch.randn_like(o))
del qkv, o

# ---- loss + optimiser --------------------------------------------------------------------------------------------
Tt, V = 4096, 151936
lg = bf(Tt, V)
tgt = torch.randint(0, V, (Tt,), device=dev)
stats = torch.empty(Tt, 3, device=dev)
lse = torch.zeros(Tt, device=dev)
N.call("stb_ce_stats", lg.data_ptr(), tgt.data_ptr(), stats.data_ptr(), Tt, V, V, 0, N.stream_ptr())
N.call("stb_ce_grad", lg.data_ptr(), tgt.data_ptr(), lse.data_ptr(), Tt, V, V, 0, 1.0, None, -100, N.stream_ptr())
del lg
from scaletorch_b200.ops.optim import fused_adamw_step, sum_of_squares  # noqa: E402

n = 150_000_000
p = torch.zeros(n, device=dev, dtype=torch.bfloat16)
g = torch.zeros(n, device=dev)
m, vv = torch.zeros_like(g), torch.zeros_like(g)
fused_adamw_step(p, g, m, vv, lr=1e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1, step=1)
sum_of_squares(g)
torch.cuda.synchronize()
print("ok")
