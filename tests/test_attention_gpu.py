"""Flash-attention kernels (tcgen05) against the fp32 eager oracle."""
import math

import pytest
import torch

from scaletorch_b200 import ops
from scaletorch_b200.ops import attention as A

pytestmark = pytest.mark.gpu


def _bf(*shape):
    return torch.randn(*shape, device="cuda").bfloat16()


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


@pytest.mark.parametrize("B,Sq,Sk,Hq,Hkv,causal", [
    (1, 128, 128, 1, 1, True), (2, 256, 256, 4, 2, True), (1, 512, 512, 8, 2, False), (2, 384, 384, 2, 1, True),
    (1, 200, 200, 2, 2, True), (1, 130, 333, 4, 1, False), (1, 2048, 2048, 4, 2, True), (1, 128, 512, 2, 1, False),
])
def test_flash_fwd_matches_oracle(B, Sq, Sk, Hq, Hkv, causal):
    if not A.native_attention_fwd_available():
        pytest.skip("native attention forward not built")
    torch.manual_seed(0)
    q, k, v = _bf(B, Sq, Hq, 128), _bf(B, Sk, Hkv, 128), _bf(B, Sk, Hkv, 128)
    scale = 1.0 / math.sqrt(128)
    out, lse = A._flash_fwd(q, k, v, causal, scale)
    ref, ref_lse = ops.attention_ref(q, k, v, causal, scale, q_offset=(Sk - Sq) if causal else 0)
    assert _rel(out, ref) < 2e-2
    assert (lse - ref_lse).abs().max().item() < 2e-2


def test_flash_fwd_reads_packed_qkv_views():
    if not A.native_attention_fwd_available():
        pytest.skip("native attention forward not built")
    torch.manual_seed(1)
    B, S, Hq, Hkv, D = 2, 256, 4, 2, 128
    qkv = _bf(B * S, (Hq + 2 * Hkv) * D)
    v4 = qkv.view(B, S, Hq + 2 * Hkv, D)
    q, k, v = v4[:, :, :Hq], v4[:, :, Hq:Hq + Hkv], v4[:, :, Hq + Hkv:]
    out, lse = A._flash_fwd(q, k, v, True, 0.09)
    ref, ref_lse = ops.attention_ref(q, k, v, True, 0.09)
    assert _rel(out, ref) < 2e-2
    assert (lse - ref_lse).abs().max().item() < 2e-2


@pytest.mark.parametrize("fused", ["1", "0"], ids=["one-kernel", "two-kernel"])
@pytest.mark.parametrize("B,S,Hq,Hkv,causal", [(1, 128, 1, 1, True), (2, 256, 4, 2, True), (1, 384, 4, 1, False),
                                               (1, 200, 2, 2, True), (1, 1024, 4, 2, True), (2, 2048, 8, 2, True)])
def test_flash_bwd_matches_oracle(B, S, Hq, Hkv, causal, fused, monkeypatch):
    if not A.native_attention_available():
        pytest.skip("native attention backward not built")
    monkeypatch.setenv("SCALETORCH_B200_ATTN_BWD_FUSED", fused)
    assert A.fused_bwd_available() == (fused == "1")
    torch.manual_seed(2)
    q, k, v = _bf(B, S, Hq, 128).requires_grad_(), _bf(B, S, Hkv, 128).requires_grad_(), _bf(B, S, Hkv, 128).requires_grad_()
    scale = 1.0 / math.sqrt(128)
    out, _ = ops.flash_attention(q, k, v, causal, scale)
    do = _bf(B, S, Hq, 128)
    out.backward(do)
    qr, kr, vr = (t.detach().float().requires_grad_() for t in (q, k, v))
    ref, _ = ops.attention_ref(qr, kr, vr, causal, scale)
    ref.backward(do.float())
    assert _rel(out, ref) < 2e-2
    assert _rel(q.grad, qr.grad) < 3e-2
    assert _rel(k.grad, kr.grad) < 3e-2
    assert _rel(v.grad, vr.grad) < 3e-2


@pytest.mark.parametrize("fused", ["1", "0"], ids=["one-kernel", "two-kernel"])
def test_qkv_attention_block_matches_unfused(fused, monkeypatch):
    if not A.native_attention_available():
        pytest.skip("native attention backward not built")
    monkeypatch.setenv("SCALETORCH_B200_ATTN_BWD_FUSED", fused)
    torch.manual_seed(3)
    B, S, Hq, Hkv, D = 2, 256, 4, 2, 128
    qkv = _bf(B * S, (Hq + 2 * Hkv) * D).requires_grad_()
    wq = (1 + 0.1 * torch.randn(D, device="cuda")).bfloat16().requires_grad_()
    wk = (1 + 0.1 * torch.randn(D, device="cuda")).bfloat16().requires_grad_()
    cos, sin = ops.get_cos_sin(512, D, 1e6, device="cuda")
    out = ops.qkv_attention(qkv, wq, wk, cos, sin, B, S, Hq, Hkv, D)
    do = _bf(B, S, Hq * D)
    out.backward(do)
    # oracle: fp32 everything
    qr = qkv.detach().float().requires_grad_()
    wqr, wkr = wq.detach().float().requires_grad_(), wk.detach().float().requires_grad_()
    qk = ops.qknorm_rope_ref(qr, wqr, wkr, cos, sin, Hq, Hkv, D, S, 0, 1e-6)
    qk4 = qk.view(B, S, Hq + Hkv, D)
    v = qr.view(B, S, Hq + 2 * Hkv, D)[:, :, Hq + Hkv:]
    ref, _ = ops.attention_ref(qk4[:, :, :Hq], qk4[:, :, Hq:], v, True, 1 / math.sqrt(D))
    ref.reshape(B, S, Hq * D).backward(do.float())
    assert _rel(out, ref.reshape(B, S, Hq * D)) < 2e-2
    assert _rel(qkv.grad, qr.grad) < 4e-2
    assert _rel(wq.grad, wqr.grad) < 5e-2
    assert _rel(wk.grad, wkr.grad) < 5e-2
