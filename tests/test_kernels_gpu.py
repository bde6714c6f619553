"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference of the same op."""
import math

import pytest
import torch

from scaletorch_b200 import _native as N
from scaletorch_b200 import ops

pytestmark = pytest.mark.gpu


def _bf(*shape, scale=1.0, device="cuda"):
    return (torch.randn(*shape, device=device, dtype=torch.float32) * scale).to(torch.bfloat16)


def _rel_err(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def test_native_library_loaded():
    assert N.load(required=True) is not None
    assert N.available()


@pytest.mark.parametrize("M,Nn,K", [(128, 128, 64), (256, 256, 128), (512, 384, 1024), (4096, 3072, 1024),
                                    (300, 200, 136), (1000, 1544, 72)])
@pytest.mark.parametrize("bn", [0, 128, 256])
def test_gemm_nt(M, Nn, K, bn):
    torch.manual_seed(0)
    a, b = _bf(M, K), _bf(Nn, K)
    c = ops.gemm(a, b, force_bn=bn)
    ref = a.float() @ b.float().t()
    assert _rel_err(c, ref) < 5e-3


@pytest.mark.parametrize("M,Nn,K", [(256, 256, 128), (512, 1024, 384), (300, 200, 136), (4096, 1024, 3072)])
def test_gemm_dgrad_layout(M, Nn, K):
    """A K-major, B MN-major:  C = A · B  with B given as [K, N]."""
    torch.manual_seed(1)
    a, b = _bf(M, K), _bf(K, Nn)
    c = ops.gemm(a, b, b_mn=True)
    ref = a.float() @ b.float()
    assert _rel_err(c, ref) < 5e-3


@pytest.mark.parametrize("M,Nn,K", [(256, 256, 128), (1024, 384, 512), (200, 136, 300), (3072, 1024, 4096)])
def test_gemm_wgrad_layout(M, Nn, K):
    """A and B MN-major: C = Aᵀ · B with A [K, M], B [K, N]; fp32 output with accumulation."""
    torch.manual_seed(2)
    a, b = _bf(K, M), _bf(K, Nn)
    ref = a.float().t() @ b.float()
    c = ops.gemm(a, b, a_mn=True, b_mn=True, out_dtype=torch.float32)
    assert _rel_err(c, ref) < 5e-3
    ops.gemm(a, b, a_mn=True, b_mn=True, out=c, accumulate=True)
    assert _rel_err(c, 2 * ref) < 5e-3


@pytest.mark.parametrize("M,Nn,K", [(256, 256, 64), (512, 512, 256), (4096, 3072, 1024), (1000, 1544, 200)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True)])
@pytest.mark.parametrize("bn", [128, 256])
def test_gemm_pair_kernel(M, Nn, K, a_mn, b_mn, bn):
    """CTA-pair (cta_group::2) kernel called directly against an fp32 matmul: store, fp32 accumulate, ragged edges."""
    from scaletorch_b200.ops.gemm import gemm_pair

    torch.manual_seed(0)
    a = torch.randn((K, M) if a_mn else (M, K), device="cuda").bfloat16()
    b = torch.randn((K, Nn) if b_mn else (Nn, K), device="cuda").bfloat16()
    ref = (a.float().t() if a_mn else a.float()) @ (b.float() if b_mn else b.float().t())
    out = torch.full((M, Nn), float("nan"), device="cuda", dtype=torch.bfloat16)
    gemm_pair(a, b, out, a_mn=a_mn, b_mn=b_mn, bn=bn, splits=1)
    torch.cuda.synchronize()
    assert _rel_err(out, ref) < 5e-3
    acc = torch.ones((M, Nn), device="cuda", dtype=torch.float32)
    gemm_pair(a, b, acc, a_mn=a_mn, b_mn=b_mn, accumulate=True, bn=bn, splits=1)
    torch.cuda.synchronize()
    assert _rel_err(acc, ref + 1.0) < 5e-3


@pytest.mark.parametrize("splits", [0, 2, 3, 7])
@pytest.mark.parametrize("accumulate", [False, True])
@pytest.mark.parametrize("bn", [128, 256])
def test_gemm_pair_split_k(splits, accumulate, bn):
    """Weight-gradient shape (few tiles, long reduction): every k-split reduce-adds into the fp32 destination."""
    from scaletorch_b200.ops.gemm import gemm_pair

    torch.manual_seed(1)
    M, Nn, K = 1024, 2048, 4096 + 64          # K not a multiple of the split
    a = torch.randn(K, M, device="cuda").bfloat16()
    b = torch.randn(K, Nn, device="cuda").bfloat16()
    ref = a.float().t() @ b.float()
    out = torch.full((M, Nn), 3.0 if accumulate else float("nan"), device="cuda", dtype=torch.float32)
    gemm_pair(a, b, out, a_mn=True, b_mn=True, accumulate=accumulate, bn=bn, splits=splits)
    torch.cuda.synchronize()
    assert _rel_err(out, ref + (3.0 if accumulate else 0.0)) < 5e-3


def test_gemm_default_dispatch_matches_single_cta():
    """``ops.gemm`` (library-chosen kernel) and the single-CTA kernel (forced tile width) agree on model shapes."""
    torch.manual_seed(2)
    for (M, Nn, K) in [(4096, 4096, 1024), (4096, 1024, 3072), (384, 640, 512)]:
        a = torch.randn(M, K, device="cuda").bfloat16()
        b = torch.randn(Nn, K, device="cuda").bfloat16()
        ref = a.float() @ b.float().t()
        assert _rel_err(ops.gemm(a, b), ref) < 5e-3
        assert _rel_err(ops.gemm(a, b, force_bn=256), ref) < 5e-3
        w = torch.randn(M, Nn, device="cuda").bfloat16()          # wgrad: [K=M tokens] reduction, fp32 out
        dw = torch.zeros(Nn, K, device="cuda")
        ops.gemm(w, a, a_mn=True, b_mn=True, out=dw)
        assert _rel_err(dw, w.float().t() @ a.float()) < 5e-3


def test_gemm_strided_views():
    torch.manual_seed(3)
    big = _bf(512, 1024)
    a = big[:, 256:768]  # lda = 1024, K = 512
    b = _bf(256, 512)
    out = torch.zeros(512, 512, device="cuda", dtype=torch.bfloat16)
    ops.gemm(a, b, out=out[:, 128:384])
    assert _rel_err(out[:, 128:384], a.float() @ b.float().t()) < 5e-3
    assert out[:, :128].abs().max() == 0 and out[:, 384:].abs().max() == 0


def test_linear_autograd():
    torch.manual_seed(4)
    x = _bf(4, 96, 256).requires_grad_()
    w = _bf(384, 256, scale=0.05).requires_grad_()
    y = ops.linear(x, w)
    dy = _bf(4, 96, 384)
    y.backward(dy)
    xr = x.detach().float().requires_grad_()
    wr = w.detach().float().requires_grad_()
    yr = torch.nn.functional.linear(xr, wr)
    yr.backward(dy.float())
    assert _rel_err(y, yr) < 5e-3
    assert _rel_err(x.grad, xr.grad) < 5e-3
    assert _rel_err(w.grad, wr.grad) < 5e-3


def test_linear_main_grad_accumulation():
    torch.manual_seed(5)
    x = _bf(256, 128)
    w = _bf(192, 128, scale=0.05).requires_grad_()
    w.main_grad = torch.full((192 * 128,), 7.0, device="cuda")  # garbage: first write must overwrite
    w.grad_written = False
    for _ in range(2):
        ops.linear(x, w).backward(torch.ones(256, 192, device="cuda", dtype=torch.bfloat16))
    assert w.grad is None
    ref = 2 * torch.ones(256, 192, device="cuda").t() @ x.float()
    assert _rel_err(w.main_grad.view(192, 128), ref) < 5e-3


@pytest.mark.parametrize("H", [128, 256, 768, 1024, 2048, 4096, 5120])
@pytest.mark.parametrize("with_res", [False, True])
def test_rmsnorm(H, with_res):
    torch.manual_seed(6)
    rows = 333
    x = _bf(rows, H).requires_grad_()
    res = _bf(rows, H).requires_grad_() if with_res else None
    w = (1 + 0.1 * torch.randn(H, device="cuda")).to(torch.bfloat16).requires_grad_()
    dy, dh = _bf(rows, H), _bf(rows, H)
    if with_res:
        h, y = ops.add_rmsnorm(x, res, w, 1e-6)
        torch.autograd.backward([h, y], [dh, dy])
    else:
        y = ops.rmsnorm(x, w, 1e-6)
        y.backward(dy)
    xr = x.detach().float().requires_grad_()
    wr = w.detach().float().requires_grad_()
    rr = res.detach().float().requires_grad_() if with_res else None
    hr = xr + rr if with_res else xr
    yr = wr * (hr * torch.rsqrt(hr.pow(2).mean(-1, keepdim=True) + 1e-6))
    if with_res:
        torch.autograd.backward([hr, yr], [dh.float(), dy.float()])
        assert _rel_err(h, hr) < 5e-3
        assert _rel_err(res.grad, rr.grad) < 1e-2
    else:
        yr.backward(dy.float())
    assert _rel_err(y, yr) < 1e-2
    assert _rel_err(x.grad, xr.grad) < 1e-2
    assert _rel_err(w.grad, wr.grad) < 2e-2


@pytest.mark.parametrize("use_norm", [True, False])
def test_qknorm_rope(use_norm):
    torch.manual_seed(7)
    B, S, Hq, Hkv, D = 2, 96, 4, 2, 128
    T = B * S
    qkv = _bf(T, (Hq + 2 * Hkv) * D).requires_grad_()
    wq = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16).requires_grad_() if use_norm else None
    wk = (1 + 0.1 * torch.randn(D, device="cuda")).to(torch.bfloat16).requires_grad_() if use_norm else None
    cos, sin = ops.get_cos_sin(256, D, 1e6, device="cuda")
    out = ops.qknorm_rope(qkv, wq, wk, cos, sin, Hq, Hkv, D, S, pos_offset=3)
    g = _bf(T, (Hq + Hkv) * D)
    out.backward(g)

    qr = qkv.detach().float().requires_grad_()
    wqr = wq.detach().float().requires_grad_() if use_norm else None
    wkr = wk.detach().float().requires_grad_() if use_norm else None
    x = qr[:, : (Hq + Hkv) * D].reshape(T, Hq + Hkv, D)
    if use_norm:
        w = torch.cat([wqr.expand(Hq, D), wkr.expand(Hkv, D)], 0)
        x = w * (x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + 1e-6))
    pos = torch.arange(T, device="cuda") % S + 3
    c, s = cos[pos].unsqueeze(1), sin[pos].unsqueeze(1)
    ref = (x * c + ops.rotate_half(x) * s).reshape(T, -1)
    ref.backward(g.float())
    assert _rel_err(out, ref) < 1e-2
    assert _rel_err(qkv.grad[:, : (Hq + Hkv) * D], qr.grad[:, : (Hq + Hkv) * D]) < 1e-2
    assert qkv.grad[:, (Hq + Hkv) * D:].abs().max() == 0
    if use_norm:
        assert _rel_err(wq.grad, wqr.grad) < 2e-2
        assert _rel_err(wk.grad, wkr.grad) < 2e-2


def test_swiglu():
    torch.manual_seed(8)
    gu = _bf(3, 100, 2 * 768).requires_grad_()
    y = ops.swiglu(gu)
    dy = _bf(3, 100, 768)
    y.backward(dy)
    gr = gu.detach().float().requires_grad_()
    yr = ops.swiglu_ref(gr)
    yr.backward(dy.float())
    assert _rel_err(y, yr) < 1e-2
    assert _rel_err(gu.grad, gr.grad) < 1e-2


@pytest.mark.parametrize("V", [1000, 151936 // 8])
def test_linear_cross_entropy(V):
    torch.manual_seed(9)
    T, H = 300, 256
    x = _bf(T, H).requires_grad_()
    w = _bf(V, H, scale=0.05).requires_grad_()
    tgt = torch.randint(0, V, (T,), device="cuda")
    tgt[::17] = -100
    loss = ops.linear_cross_entropy(x, w, tgt, loss_scale=0.5)
    (loss * 2.0).backward()
    xr = x.detach().float().requires_grad_()
    wr = w.detach().float().requires_grad_()
    lr = torch.nn.functional.cross_entropy(xr @ wr.t(), tgt, ignore_index=-100) * 0.5
    (lr * 2.0).backward()
    assert abs(loss.item() - lr.item()) < 2e-2 * abs(lr.item())
    assert _rel_err(x.grad, xr.grad) < 3e-2
    assert _rel_err(w.grad, wr.grad) < 3e-2


def test_linear_cross_entropy_chunked(monkeypatch):
    """Vocabulary processed in column chunks (online statistics, per-chunk gradient / dgrad / wgrad into main_grad)."""
    monkeypatch.setenv("SCALETORCH_B200_CE_CHUNK_MB", "1")
    torch.manual_seed(9)
    T, H, V = 384, 256, 20000                      # 1 MB budget -> chunks of 1280 columns, ragged last chunk
    x = _bf(T, H).requires_grad_()
    w = torch.nn.Parameter(_bf(V, H, scale=0.05))
    w.main_grad = torch.full((V * H,), float("nan"), device="cuda")
    w.grad_written = False
    tgt = torch.randint(0, V, (T,), device="cuda")
    tgt[::13] = -100
    from scaletorch_b200.ops.cross_entropy import _chunks

    assert len(_chunks(T, V)) > 8
    loss = ops.linear_cross_entropy(x, w, tgt, loss_scale=0.5)
    (loss * 2.0).backward()
    xr = x.detach().float().requires_grad_()
    wr = w.detach().float().requires_grad_()
    lr = torch.nn.functional.cross_entropy(xr @ wr.t(), tgt, ignore_index=-100) * 0.5
    (lr * 2.0).backward()
    assert abs(loss.item() - lr.item()) < 2e-2 * abs(lr.item())
    assert _rel_err(x.grad, xr.grad) < 3e-2
    assert w.grad is None and w.grad_written
    assert _rel_err(w.main_grad.view(V, H), wr.grad) < 3e-2
    # second backward accumulates into main_grad
    loss2 = ops.linear_cross_entropy(x, w, tgt, loss_scale=0.5)
    (loss2 * 2.0).backward()
    assert _rel_err(w.main_grad.view(V, H), 2 * wr.grad) < 3e-2


def test_embedding():
    torch.manual_seed(10)
    V, H = 1000, 256
    w = _bf(V, H).requires_grad_()
    ids = torch.randint(0, V, (4, 77), device="cuda")
    out = ops.embedding(ids, w)
    dy = _bf(4, 77, H)
    out.backward(dy)
    wr = w.detach().float().requires_grad_()
    outr = torch.nn.functional.embedding(ids, wr)
    outr.backward(dy.float())
    assert torch.equal(out.float(), outr)
    assert _rel_err(w.grad, wr.grad) < 1e-2
    # vocab shard: ids outside [start, start+V) give zero rows
    out2 = ops.embedding(ids + 500, w, vocab_start=500)
    assert torch.equal(out2, out)
    out3 = ops.embedding(ids, w[:500], vocab_start=0)
    assert (out3[ids >= 500] == 0).all()


def test_adamw_and_gradnorm():
    from scaletorch_b200.ops.optim import fused_adamw_step, grad_clip_coef

    torch.manual_seed(11)
    n = 1_000_003
    p = _bf(n)
    g = torch.randn(n, device="cuda")
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    pr = p.float().clone().requires_grad_()
    opt = torch.optim.AdamW([pr], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    coef, norm = grad_clip_coef(g, max_norm=1.0)
    assert abs(norm.item() - g.norm().item()) < 1e-3 * g.norm().item()
    for step in (1, 2):
        pr.grad = g * min(1.0, 1.0 / (g.norm().item() + 1e-6))
        opt.step()
        fused_adamw_step(p, g, m, v, lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1, step=step, clip_coef=coef)
    assert _rel_err(p, pr.detach()) < 1e-2
    assert _rel_err(m, opt.state[pr]["exp_avg"]) < 1e-3
    assert _rel_err(v, opt.state[pr]["exp_avg_sq"]) < 1e-3


@pytest.mark.parametrize("E,N_,K,counts", [(4, 256, 128, [100, 0, 300, 129]), (8, 1536, 2048, [512, 37, 0, 1000, 128, 1, 700, 256]),
                                            (3, 128, 64, [0, 0, 5])])
def test_grouped_gemm_modes(E, N_, K, counts):
    """Ragged grouped GEMM (fprop / dgrad / wgrad) vs a per-expert fp32 loop."""
    from scaletorch_b200.ops.grouped import _GroupedLinear, native_grouped_available
    from scaletorch_b200.ops.grouped_native import grouped_linear_native

    assert native_grouped_available()
    torch.manual_seed(12)
    R = sum(counts)
    tpe = torch.tensor(counts, device="cuda")
    x = _bf(R, K).requires_grad_()
    w = _bf(E, N_, K, scale=0.05).requires_grad_()
    y = grouped_linear_native(x, w, tpe)
    dy = _bf(R, N_)
    y.backward(dy)
    xr, wr = x.detach().float().requires_grad_(), w.detach().float().requires_grad_()
    outs, off = [], 0
    for e, c in enumerate(counts):
        outs.append(xr[off:off + c] @ wr[e].t())
        off += c
    yr = torch.cat(outs)
    yr.backward(dy.float())
    assert _rel_err(y, yr) < 1e-2
    assert _rel_err(x.grad, xr.grad) < 1e-2
    assert _rel_err(w.grad, wr.grad) < 1e-2
    # accumulation into main_grad (fp32) with experts that received no token
    w2 = w.detach().clone().requires_grad_()
    w2.main_grad = torch.full((E * N_ * K,), 3.0, device="cuda")
    w2.grad_written = False
    for _ in range(2):
        grouped_linear_native(x.detach(), w2, tpe).backward(dy)
    assert _rel_err(w2.main_grad.view(E, N_, K), 2 * wr.grad) < 1e-2


def test_moe_layer_native_matches_oracle():
    """Qwen3-MoE layer (router → permute → grouped SwiGLU experts → combine) native vs PyTorch oracle path."""
    import os
    from types import SimpleNamespace

    from scaletorch_b200.models.model_qwen3_moe import MoELayer
    from scaletorch_b200.parallel.context import ParallelContext

    cfg = SimpleNamespace(num_experts=8, num_experts_per_tok=2, hidden_size=256, moe_intermediate_size=128,
                          norm_topk_prob=True, router_aux_loss_coef=0.001)
    torch.manual_seed(13)
    layer = MoELayer(cfg, ParallelContext()).cuda().bfloat16()
    x = _bf(700, 256).requires_grad_()
    y = layer(x)
    g = _bf(700, 256)
    y.backward(g)
    grads = {n: p.grad.clone() for n, p in layer.named_parameters()}
    xg = x.grad.clone()
    layer.zero_grad()
    x.grad = None
    os.environ["SCALETORCH_B200_FORCE_TORCH"] = "1"
    try:
        y2 = layer(x)
        y2.backward(g)
    finally:
        os.environ["SCALETORCH_B200_FORCE_TORCH"] = "0"
    assert _rel_err(y, y2) < 3e-2
    assert _rel_err(xg, x.grad) < 5e-2
    for n, p in layer.named_parameters():
        assert _rel_err(grads[n], p.grad) < 6e-2, n


@pytest.mark.parametrize("T,E,k", [(777, 128, 8), (64, 8, 2), (1000, 60, 4), (33, 256, 16)])
@pytest.mark.parametrize("coef", [0.0, 0.01])
def test_moe_router_kernels(T, E, k, coef):
    """Routing kernel (top-k, gate weights, counts, positions, aux loss) and its backward vs the PyTorch oracle."""
    from scaletorch_b200.ops.moe_routing import route, route_ref

    torch.manual_seed(3)
    # distinct bf16 logits per row so that top-k has no ties to break differently
    base = torch.stack([torch.randperm(E, device="cuda").float() for _ in range(T)]) / 8.0 - E / 16.0
    logits = base.bfloat16().requires_grad_()
    w, idx, pos, counts, aux = route(logits, k, coef, True)
    assert idx.dtype == torch.int32 and pos.dtype == torch.int32
    lr = logits.detach().clone().requires_grad_()
    w_r, idx_r, pos_r, counts_r, aux_r = route_ref(lr, k, coef, True)
    assert torch.equal(idx.long(), idx_r)
    assert torch.equal(counts.long(), counts_r)
    assert _rel_err(w, w_r) < 1e-5
    assert abs(float(aux) - float(aux_r)) <= 2e-3 * max(1e-3, abs(float(aux_r))), (float(aux), float(aux_r))
    # positions: a bijection onto [0, count) inside every expert
    flat, p = idx.reshape(-1).long(), pos.reshape(-1).long()
    key = flat * (T * k) + p
    assert key.unique().numel() == T * k
    assert bool((p < counts.long()[flat]).all()) and bool((p >= 0).all())
    gw = torch.randn(T, k, device="cuda")
    (w * gw).sum().add(aux * 3.0).backward()
    (w_r * gw).sum().add(aux_r * 3.0).backward()
    assert _rel_err(logits.grad, lr.grad) < 2e-2, _rel_err(logits.grad, lr.grad)


@pytest.mark.parametrize("ep,E,cap_factor", [(4, 32, 4.0), (2, 8, 1.1), (8, 128, 8.0)])
def test_ep_plan_kernel_matches_tensor_plan(ep, E, cap_factor):
    """``stb_ep_plan`` against the pure-tensor plan: same segments, same owner ranks, rows a bijection per expert."""
    from scaletorch_b200.parallel.expert_parallel import fused_ep  # registers the kernel signatures

    torch.manual_seed(11)
    T, k = 500, 2
    e_local = E // ep
    idx = [torch.stack([torch.randperm(E, device="cuda")[:k] for _ in range(T)]) for _ in range(ep)]
    table = torch.stack([torch.bincount(i.reshape(-1), minlength=E) for i in idx]).to(torch.int32)
    rows_cap = (int(cap_factor * T * k) + e_local * 127 + 127) // 128 * 128
    for r in range(ep):
        flat = idx[r].reshape(-1)
        row_t, dest_t, drop_t, offs_t, tpe_t = fused_ep.ep_plan(table.long(), flat, table[r].long(), r, e_local, rows_cap)
        order = torch.argsort(flat, stable=True)
        inv = torch.empty_like(order)
        inv[order] = torch.arange(flat.numel(), device="cuda")
        c = table[r].long()
        pos = (inv - (torch.cumsum(c, 0) - c)[flat]).to(torch.int32)
        n_tiles = rows_cap // 128
        row = torch.empty(T * k, dtype=torch.int32, device="cuda")
        dest = torch.empty_like(row)
        dropped = torch.zeros(1, dtype=torch.int32, device="cuda")
        offs = torch.empty(e_local + 1, dtype=torch.int32, device="cuda")
        tpe = torch.empty(e_local, dtype=torch.int32, device="cuda")
        tile_e = torch.empty(n_tiles, dtype=torch.int32, device="cuda")
        N.call("stb_ep_plan", table.data_ptr(), flat.to(torch.int32).data_ptr(), pos.data_ptr(), T * k, E, ep, r, rows_cap,
               row.data_ptr(), dest.data_ptr(), dropped.data_ptr(), offs.data_ptr(), tpe.data_ptr(), tile_e.data_ptr(),
               n_tiles, N.stream_ptr())
        torch.cuda.synchronize()
        assert torch.equal(row, row_t) and torch.equal(dest, dest_t)
        assert int(dropped) == int(drop_t)
        assert torch.equal(offs.long(), offs_t.long()) and torch.equal(tpe.long(), tpe_t.long())
        lay = fused_ep.GroupedLayout.from_padded_offsets(offs_t, rows_cap, e_local)
        assert torch.equal(tile_e, lay.tile_expert)
