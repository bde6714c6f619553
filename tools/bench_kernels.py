"""Micro-benchmarks of the sm_100a kernels against the library path they replace.

CUDA-event timing, warm-up, L2 flushed between iterations (a 256 MB buffer is written), results
as JSON lines in ``gpurun_out/kernels.jsonl``.  Usage: ``python tools/bench_kernels.py [gemm|mem|attn|all]``.
"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaletorch_b200 import ops  # noqa: E402

PEAKS = {}
try:
    PEAKS = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
except Exception:
    pass
_flush = None


def timeit(fn, iters=10, warmup=3, flush=True):
    global _flush
    if _flush is None:
        _flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush:
            _flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def emit(rec, out):
    print(json.dumps(rec), flush=True)
    out.write(json.dumps(rec) + "\n")
    out.flush()


def bench_gemm(out):
    peak = PEAKS.get("bf16_tflops", 1590.0)
    shapes = [  # (name, M, N, K)  Qwen3-0.6B at 4096 tokens, 8B TP2 at 4096 tokens, square
        ("0.6B qkv", 4096, 4096, 1024), ("0.6B o", 4096, 1024, 2048), ("0.6B gate_up", 4096, 6144, 1024),
        ("0.6B down", 4096, 1024, 3072), ("0.6B lm_head", 4096, 151936, 1024),
        ("8B-tp2 qkv", 4096, 3072, 4096), ("8B-tp2 gate_up", 4096, 12288, 4096), ("8B-tp2 down", 4096, 4096, 6144),
        ("square 4096", 4096, 4096, 4096), ("square 8192", 8192, 8192, 8192),
    ]
    for name, M, N, K in shapes:
        for mode in ("fprop", "dgrad", "wgrad"):
            if mode == "fprop":
                a = torch.randn(M, K, device="cuda").bfloat16(); b = torch.randn(N, K, device="cuda").bfloat16()
                ours = lambda: ops.gemm(a, b)
                lib = lambda: torch.matmul(a, b.t())
                flops = 2.0 * M * N * K
            elif mode == "dgrad":  # dx[M,K] = dy[M,N] @ W[N,K]
                a = torch.randn(M, N, device="cuda").bfloat16(); b = torch.randn(N, K, device="cuda").bfloat16()
                ours = lambda: ops.gemm(a, b, b_mn=True)
                lib = lambda: torch.matmul(a, b)
                flops = 2.0 * M * N * K
            else:  # dW[N,K] = dy[M,N]^T @ x[M,K]  (fp32 out)
                a = torch.randn(M, N, device="cuda").bfloat16(); b = torch.randn(M, K, device="cuda").bfloat16()
                o = torch.empty(N, K, device="cuda", dtype=torch.float32)
                ours = lambda: ops.gemm(a, b, a_mn=True, b_mn=True, out=o)
                lib = lambda: torch.matmul(a.t(), b)
                flops = 2.0 * M * N * K
            t_o = timeit(ours)
            t_l = timeit(lib)
            emit({"kernel": "gemm", "shape": name, "mode": mode, "M": M, "N": N, "K": K,
                  "ours_ms": round(t_o, 4), "cublas_ms": round(t_l, 4),
                  "ours_tflops": round(flops / t_o / 1e9, 1), "cublas_tflops": round(flops / t_l / 1e9, 1),
                  "frac_of_measured_peak": round(flops / t_o / 1e9 / peak, 3), "vs_cublas": round(t_l / t_o, 3)}, out)
            del a, b


def bench_mem(out):
    bw = PEAKS.get("hbm_gbs", 6650.0)
    T, H, I = 8192, 4096, 12288
    x = torch.randn(T, H, device="cuda").bfloat16(); r = torch.randn(T, H, device="cuda").bfloat16()
    w = torch.ones(H, device="cuda").bfloat16()
    t = timeit(lambda: ops.add_rmsnorm(x, r, w, 1e-6))
    emit({"kernel": "add_rmsnorm_fwd", "T": T, "H": H, "ms": round(t, 4), "GBps": round(4 * T * H * 2 / t / 1e6, 1),
          "frac_of_measured_hbm": round(4 * T * H * 2 / t / 1e6 / bw, 3)}, out)
    t = timeit(lambda: torch.nn.functional.rms_norm(x + r, (H,), w.float(), 1e-6))
    emit({"kernel": "torch add+rms_norm", "T": T, "H": H, "ms": round(t, 4)}, out)
    # Qwen3-0.6B shapes: RMSNorm backward at H = 1024 and fused QK-norm + RoPE forward / backward (16 q + 8 kv heads)
    T2, H2 = 4096, 1024
    x2 = torch.randn(T2, H2, device="cuda").bfloat16().requires_grad_()
    r2 = torch.randn(T2, H2, device="cuda").bfloat16().requires_grad_()
    w2 = torch.ones(H2, device="cuda").bfloat16().requires_grad_()
    hh, yy = ops.add_rmsnorm(x2, r2, w2, 1e-6)
    gy, gh = torch.randn_like(yy), torch.randn_like(hh)
    t = timeit(lambda: torch.autograd.grad((yy, hh), (x2, r2, w2), (gy, gh), retain_graph=True))
    emit({"kernel": "add_rmsnorm_bwd (+dw reduce)", "T": T2, "H": H2, "ms": round(t, 4), "GBps": round(4 * T2 * H2 * 2 / t / 1e6, 1),
          "frac_of_measured_hbm": round(4 * T2 * H2 * 2 / t / 1e6 / bw, 3)}, out)
    from scaletorch_b200.ops import attention as A
    nq, nkv, D = 16, 8, 128
    qkv = torch.randn(T2, (nq + 2 * nkv) * D, device="cuda").bfloat16()
    wq = torch.ones(D, device="cuda").bfloat16(); wk = torch.ones(D, device="cuda").bfloat16()
    cos, sin = ops.get_cos_sin(4096, D, 1e6, device="cuda")
    qk = torch.empty(T2, (nq + nkv) * D, device="cuda", dtype=torch.bfloat16)
    rstd = torch.empty(T2 * (nq + nkv), device="cuda")
    from scaletorch_b200 import _native as N
    f = lambda: N.call("stb_qknorm_rope_fwd", qkv.data_ptr(), wq.data_ptr(), wk.data_ptr(), cos.data_ptr(), sin.data_ptr(), None,
                       qk.data_ptr(), rstd.data_ptr(), T2, 2048, nq, nkv, D, 0, 1, 1e-6, N.stream_ptr())
    t = timeit(f)
    nb = 2 * T2 * (nq + nkv) * D * 2
    emit({"kernel": "qknorm_rope_fwd", "T": T2, "heads": nq + nkv, "ms": round(t, 4), "GBps": round(nb / t / 1e6, 1), "frac_of_measured_hbm": round(nb / t / 1e6 / bw, 3)}, out)
    dqk = torch.randn_like(qk); dqkv = torch.empty_like(qkv); dwq = torch.zeros(D, device="cuda"); dwk = torch.zeros(D, device="cuda")
    b = lambda: N.call("stb_qknorm_rope_bwd", dqk.data_ptr(), qkv.data_ptr(), wq.data_ptr(), wk.data_ptr(), cos.data_ptr(), sin.data_ptr(),
                       None, rstd.data_ptr(), dqkv.data_ptr(), dwq.data_ptr(), dwk.data_ptr(), T2, 2048, nq, nkv, D, 0, 1, N.stream_ptr())
    t = timeit(b)
    nb = 3 * T2 * (nq + nkv) * D * 2
    emit({"kernel": "qknorm_rope_bwd", "T": T2, "heads": nq + nkv, "ms": round(t, 4), "GBps": round(nb / t / 1e6, 1), "frac_of_measured_hbm": round(nb / t / 1e6 / bw, 3)}, out)
    gu = torch.randn(T, 2 * I, device="cuda").bfloat16()
    t = timeit(lambda: ops.swiglu(gu))
    emit({"kernel": "swiglu_fwd", "T": T, "I": I, "ms": round(t, 4), "GBps": round(3 * T * I * 2 / t / 1e6, 1),
          "frac_of_measured_hbm": round(3 * T * I * 2 / t / 1e6 / bw, 3)}, out)
    n = 600_000_000
    p = torch.zeros(n, device="cuda", dtype=torch.bfloat16); g = torch.zeros(n, device="cuda"); m = torch.zeros_like(g); v = torch.zeros_like(g)
    from scaletorch_b200.ops.optim import fused_adamw_step, sum_of_squares
    t = timeit(lambda: fused_adamw_step(p, g, m, v, lr=1e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1, step=1), flush=False)
    emit({"kernel": "adamw", "n": n, "ms": round(t, 4), "GBps": round(n * 24 / t / 1e6, 1), "frac_of_measured_hbm": round(n * 24 / t / 1e6 / bw, 3)}, out)
    t = timeit(lambda: sum_of_squares(g), flush=False)
    emit({"kernel": "sumsq", "n": n, "ms": round(t, 4), "GBps": round(n * 4 / t / 1e6, 1), "frac_of_measured_hbm": round(n * 4 / t / 1e6 / bw, 3)}, out)
    del p, g, m, v
    Tt, V = 4096, 151936
    logits = torch.randn(Tt, V, device="cuda").bfloat16()
    tgt = torch.randint(0, V, (Tt,), device="cuda")
    stats = torch.empty(Tt, 3, device="cuda")
    from scaletorch_b200 import _native as N
    t = timeit(lambda: N.call("stb_ce_stats", logits.data_ptr(), tgt.data_ptr(), stats.data_ptr(), Tt, V, V, 0, N.stream_ptr()))
    emit({"kernel": "ce_stats", "T": Tt, "V": V, "ms": round(t, 4), "GBps": round(Tt * V * 2 / t / 1e6, 1), "frac_of_measured_hbm": round(Tt * V * 2 / t / 1e6 / bw, 3)}, out)
    lse = torch.zeros(Tt, device="cuda")
    t = timeit(lambda: N.call("stb_ce_grad", logits.data_ptr(), tgt.data_ptr(), lse.data_ptr(), Tt, V, V, 0, 1.0, None, -100, N.stream_ptr()))
    emit({"kernel": "ce_grad", "T": Tt, "V": V, "ms": round(t, 4), "GBps": round(2 * Tt * V * 2 / t / 1e6, 1), "frac_of_measured_hbm": round(2 * Tt * V * 2 / t / 1e6 / bw, 3)}, out)


def bench_attn(out):
    import torch.nn.functional as F
    peak = PEAKS.get("bf16_tflops", 1590.0)
    for (B, S, Hq, Hkv) in [(2, 2048, 16, 8), (1, 4096, 32, 8), (1, 8192, 32, 8)]:
        D = 128
        q = torch.randn(B, S, Hq, D, device="cuda").bfloat16().requires_grad_()
        k = torch.randn(B, S, Hkv, D, device="cuda").bfloat16().requires_grad_()
        v = torch.randn(B, S, Hkv, D, device="cuda").bfloat16().requires_grad_()
        flops = 4.0 * B * Hq * S * S * D / 2  # causal
        rec = {"kernel": "attention", "B": B, "S": S, "Hq": Hq, "Hkv": Hkv}
        if ops.native_attention_available():
            t = timeit(lambda: ops.flash_attention(q, k, v, True))
            rec.update(ours_fwd_ms=round(t, 4), ours_fwd_tflops=round(flops / t / 1e9, 1), frac_of_measured_peak=round(flops / t / 1e9 / peak, 3))
            o, _ = ops.flash_attention(q, k, v, True)
            do = torch.randn_like(o)
            t = timeit(lambda: torch.autograd.grad(o, (q, k, v), do, retain_graph=True))
            rec.update(ours_bwd_ms=round(t, 4), ours_bwd_tflops=round(2.5 * flops / t / 1e9, 1))
        qt, kt, vt = q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2)
        t = timeit(lambda: F.scaled_dot_product_attention(qt, kt, vt, is_causal=True, enable_gqa=True))
        rec.update(sdpa_fwd_ms=round(t, 4), sdpa_fwd_tflops=round(flops / t / 1e9, 1))
        o2 = F.scaled_dot_product_attention(qt, kt, vt, is_causal=True, enable_gqa=True)
        do2 = torch.randn_like(o2)
        t = timeit(lambda: torch.autograd.grad(o2, (q, k, v), do2, retain_graph=True))
        rec.update(sdpa_bwd_ms=round(t, 4), sdpa_bwd_tflops=round(2.5 * flops / t / 1e9, 1))
        emit(rec, out)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/kernels.jsonl", "a") as out:
        if what in ("gemm", "all"):
            bench_gemm(out)
        if what in ("mem", "all"):
            bench_mem(out)
        if what in ("attn", "all"):
            bench_attn(out)
