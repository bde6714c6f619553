"""Kernels written after the round's GPU budget was spent: compiled, never run.  Skipped unless
SCALETORCH_B200_EXPERIMENTAL=1 so an unvalidated kernel can never turn the suite red; first thing to run next round:

    SCALETORCH_B200_EXPERIMENTAL=1 timeout 120 python -m pytest tests/test_experimental_gpu.py -x -q -m gpu
"""
import os

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(os.environ.get("SCALETORCH_B200_EXPERIMENTAL", "0") != "1",
                                                  reason="experimental kernels: set SCALETORCH_B200_EXPERIMENTAL=1")]


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


@pytest.mark.parametrize("M,Nn,K", [(256, 256, 64), (512, 512, 256), (4096, 3072, 1024), (1000, 1544, 200)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True)])
def test_gemm_2cta(M, Nn, K, a_mn, b_mn):
    from scaletorch_b200.ops.gemm import _gemm_2cta

    torch.manual_seed(0)
    a = torch.randn((K, M) if a_mn else (M, K), device="cuda").bfloat16()
    b = torch.randn((K, Nn) if b_mn else (Nn, K), device="cuda").bfloat16()
    ref = (a.float().t() if a_mn else a.float()) @ (b.float() if b_mn else b.float().t())
    out = torch.full((M, Nn), float("nan"), device="cuda", dtype=torch.bfloat16)
    assert _gemm_2cta(a, b, out, M, Nn, K, a_mn, b_mn, False, force=True)
    torch.cuda.synchronize()
    assert _rel(out, ref) < 5e-3
    acc = torch.ones((M, Nn), device="cuda", dtype=torch.float32)
    assert _gemm_2cta(a, b, acc, M, Nn, K, a_mn, b_mn, True, force=True)
    torch.cuda.synchronize()
    assert _rel(acc, ref + 1.0) < 5e-3
