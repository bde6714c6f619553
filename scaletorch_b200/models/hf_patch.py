"""Monkey-patches Hugging Face ``transformers`` Qwen3 / Qwen3-MoE modules with this framework's fused ops.

B200 counterpart of the reference's ``scaletorch/models/npu_patch.py:71-240``, which swaps the same HF modules for
Ascend fused ops (``npu_rms_norm``, ``npu_swiglu`` on ``cat(gate, up)``, grouped matmul with permute / un-permute for the
MoE block).  Use it to run a stock ``AutoModelForCausalLM`` (fine-tuning scripts, evaluation harnesses) on the sm_100a
kernels without switching to the framework's own model classes:

    from scaletorch_b200.models.hf_patch import apply_b200_patch
    apply_b200_patch()                       # before or after the model is built: forwards are looked up per call
    model = AutoModelForCausalLM.from_pretrained(...)

Patched: ``Qwen3RMSNorm`` / ``Qwen3MoeRMSNorm`` → fused RMSNorm kernel; ``Qwen3MLP`` / ``Qwen3MoeMLP`` → tcgen05 GEMMs +
fused SwiGLU; ``Qwen3MoeExperts`` → expert-sorted rows through the device-scheduled grouped GEMM (no per-expert Python
loop, no ``nonzero`` host syncs).  Rotary embedding and attention keep HF's implementation (HF applies RoPE to separate
q / k tensors after its own projections; the fused QK-norm + RoPE kernel wants the packed qkv of this framework's
models).  On CPU tensors every op falls back to its PyTorch oracle, so patched and unpatched models agree there too
(``tests/test_hf_patch_cpu.py``).
"""
from __future__ import annotations

import torch

from .. import ops
from ..ops.grouped import expert_counts, grouped_mlp

_originals: dict = {}


def _rmsnorm_forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
    shape = hidden_states.shape
    y = ops.rmsnorm(hidden_states.reshape(-1, shape[-1]), self.weight, self.variance_epsilon)
    return y.reshape(shape)


def _mlp_forward(self, x: torch.Tensor) -> torch.Tensor:
    shape = x.shape
    x2 = x.reshape(-1, shape[-1])
    gu = torch.cat([ops.linear(x2, self.gate_proj.weight), ops.linear(x2, self.up_proj.weight)], dim=-1)
    y = ops.linear(ops.swiglu(gu), self.down_proj.weight)
    return y.reshape(*shape[:-1], y.shape[-1])


def _experts_forward(self, hidden_states: torch.Tensor, top_k_index: torch.Tensor, top_k_weights: torch.Tensor) -> torch.Tensor:
    """``hidden_states [T, H]``, ``top_k_index / top_k_weights [T, k]`` → ``[T, H]``: sort the (token, k) pairs by expert,
    run the grouped SwiGLU MLP over the sorted rows, weight and scatter-add back."""
    T, H = hidden_states.shape
    k = top_k_index.shape[1]
    flat = top_k_index.reshape(-1)
    order = torch.argsort(flat, stable=True)
    token_of = torch.div(order, k, rounding_mode="floor")
    xs = hidden_states.index_select(0, token_of)
    counts = expert_counts(flat, self.num_experts)
    ys = grouped_mlp(xs, counts, self.gate_up_proj, self.down_proj)
    w = top_k_weights.reshape(-1).index_select(0, order).to(ys.dtype)
    out = torch.zeros((T, H), dtype=ys.dtype, device=ys.device)
    return out.index_add(0, token_of, ys * w.unsqueeze(-1)).to(hidden_states.dtype)


def _targets():
    out = []
    try:
        from transformers.models.qwen3 import modeling_qwen3 as q3

        out += [(q3.Qwen3RMSNorm, _rmsnorm_forward), (q3.Qwen3MLP, _mlp_forward)]
    except Exception:  # pragma: no cover - transformers without Qwen3
        pass
    try:
        from transformers.models.qwen3_moe import modeling_qwen3_moe as q3m

        out += [(q3m.Qwen3MoeRMSNorm, _rmsnorm_forward), (q3m.Qwen3MoeMLP, _mlp_forward)]
        if hasattr(q3m, "Qwen3MoeExperts"):
            out.append((q3m.Qwen3MoeExperts, _experts_forward))
    except Exception:  # pragma: no cover
        pass
    return out


def apply_b200_patch() -> list[str]:
    """Installs the fused forwards; returns the names of the patched classes.  Idempotent."""
    patched = []
    for cls, fwd in _targets():
        if cls not in _originals:
            _originals[cls] = cls.forward
        cls.forward = fwd
        patched.append(cls.__name__)
    return patched


def remove_b200_patch() -> None:
    """Undo ``apply_b200_patch``: restore the original ``forward`` of every patched HuggingFace class."""
    for cls, fwd in _originals.items():
        cls.forward = fwd
    _originals.clear()
