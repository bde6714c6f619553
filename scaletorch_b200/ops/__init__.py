"""Fused ops.  Every op has two implementations behind one function:

* the hand-written sm_100a kernel from ``libscaletorch_b200.so`` (mandatory for CUDA tensors --
  a missing library raises instead of silently falling back), and
* a plain PyTorch fp32-statistics implementation that serves as the numerical oracle in the tests
  and as the CPU / gloo execution path.
"""
from .activation import swiglu, swiglu_ref
from .attention import attention_ref, flash_attention, native_attention_available, qkv_attention
from .cross_entropy import cross_entropy_ref, linear_cross_entropy
from .embedding import embedding, embedding_ref
from .gemm import gemm, linear
from .norm import (
    add_rmsnorm,
    apply_rotary_pos_emb,
    get_cos_sin,
    qknorm_rope,
    qknorm_rope_ref,
    rmsnorm,
    rmsnorm_ref,
    rotate_half,
)

__all__ = [
    "add_rmsnorm", "apply_rotary_pos_emb", "attention_ref", "cross_entropy_ref", "embedding",
    "embedding_ref", "flash_attention", "gemm", "get_cos_sin", "linear", "linear_cross_entropy",
    "native_attention_available", "qknorm_rope", "qknorm_rope_ref", "qkv_attention", "rmsnorm",
    "rmsnorm_ref", "rotate_half", "swiglu", "swiglu_ref",
]
