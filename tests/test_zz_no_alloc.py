"""`stb_rmsnorm_bwd_noalloc` -- the RMSNorm backward for callers that cannot allocate device memory.

Who needs it: programs that drive `libscaletorch_b200.so` through its C ABI from a process that has
already given all of the GPU's memory to its own arena (so there is no allocator left to ask for a
scratch buffer), or from code where the device-wide stall of a `cudaMalloc` is not tolerable.  The
library keeps the scratch for them in a fixed `__device__` array.

The test talks to the shared library directly through ctypes, exactly like such a caller would: it runs
the ordinary entry point (scratch passed in) and the no-allocation entry point on the same inputs and
requires the same `dx` / `dw`, twice in a row with different data -- the second call finds the first
call's partial sums still sitting in the shared array and must not be disturbed by them.  It also
checks that an oversized shape is refused with a message instead of being run.

This file sorts last on purpose: a `pytest -x` run has executed every other test before it gets here.
It does not depend on them -- it passes in a fresh process too.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from scaletorch_b200 import _native

    lib = ctypes.CDLL(str(_native.lib_path()))
    P, I, L = ctypes.c_void_p, ctypes.c_int, ctypes.c_long
    lib.stb_last_error.restype = ctypes.c_char_p
    lib.stb_rmsnorm_bwd_scratch_elems.restype = L
    lib.stb_rmsnorm_bwd_scratch_elems.argtypes = [I, I]
    lib.stb_rmsnorm_bwd_noalloc_max_elems.restype = L
    lib.stb_rmsnorm_bwd_noalloc_max_elems.argtypes = []
    lib.stb_rmsnorm_bwd.restype = I
    lib.stb_rmsnorm_bwd.argtypes = [P, P, P, P, P, P, P, I, P, I, I, P]
    lib.stb_rmsnorm_bwd_noalloc.restype = I
    lib.stb_rmsnorm_bwd_noalloc.argtypes = [P, P, P, P, P, P, P, I, I, I, P]
    return lib


def _inputs(rows, H, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    dy = torch.randn(rows, H, device="cuda", generator=g).bfloat16()
    h = torch.randn(rows, H, device="cuda", generator=g).bfloat16()
    dres = torch.randn(rows, H, device="cuda", generator=g).bfloat16()
    w = (1 + 0.1 * torch.randn(H, device="cuda", generator=g)).bfloat16()
    rstd = torch.rsqrt(h.float().pow(2).mean(-1) + 1e-6).contiguous()
    return dy, h, w, rstd, dres


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def test_noalloc_matches_regular_entry_point_twice():
    lib = _lib()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for call, (rows, H) in enumerate([(777, 2048), (4096, 1024)]):
        dy, h, w, rstd, dres = _inputs(rows, H, seed=100 + call)
        # ordinary entry point: the caller brings the scratch
        n = lib.stb_rmsnorm_bwd_scratch_elems(rows, H)
        scratch = torch.empty(n, device="cuda", dtype=torch.float32)
        dx_ref = torch.empty_like(h)
        dw_ref = torch.empty(H, device="cuda", dtype=torch.float32)
        rc = lib.stb_rmsnorm_bwd(dy.data_ptr(), h.data_ptr(), w.data_ptr(), rstd.data_ptr(), dres.data_ptr(),
                                 dx_ref.data_ptr(), dw_ref.data_ptr(), 0, scratch.data_ptr(), rows, H, stream)
        assert rc == 0, lib.stb_last_error()
        # no-allocation entry point: outputs are pre-made by the caller, nothing else is passed
        dx = torch.empty_like(h)
        dw = torch.empty(H, device="cuda", dtype=torch.float32)
        before = torch.cuda.memory_allocated()
        rc = lib.stb_rmsnorm_bwd_noalloc(dy.data_ptr(), h.data_ptr(), w.data_ptr(), rstd.data_ptr(), dres.data_ptr(),
                                         dx.data_ptr(), dw.data_ptr(), 0, rows, H, stream)
        assert rc == 0, lib.stb_last_error()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        assert _rel(dx, dx_ref) < 1e-2
        assert _rel(dw, dw_ref) < 2e-2
        # and both agree with a plain PyTorch fp32 formula for the weight gradient
        xn = h.float() * rstd[:, None]
        assert _rel(dw, (dy.float() * xn).sum(0)) < 2e-2


def test_noalloc_refuses_shapes_beyond_its_scratch():
    lib = _lib()
    rows, H = 100000, 8192
    assert lib.stb_rmsnorm_bwd_scratch_elems(rows, H) > lib.stb_rmsnorm_bwd_noalloc_max_elems()
    rc = lib.stb_rmsnorm_bwd_noalloc(None, None, None, None, None, None, None, 0, rows, H, None)
    assert rc != 0
    msg = lib.stb_last_error().decode()
    assert "scratch" in msg and "stb_rmsnorm_bwd" in msg
