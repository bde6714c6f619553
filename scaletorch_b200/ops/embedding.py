"""Token embedding gather / scatter-add (vocab-shard aware).

Reference: ``Qwen3Embedding`` (``scaletorch/models/model_qwen3.py:90-110``) and the masked
lookup of ``VocabParallelEmbedding`` (``scaletorch/parallel/tensor_parallel/
tensor_parallel.py:464-507``): ids outside ``[vocab_start, vocab_start + V_local)`` yield a zero
row, which the caller sums across TP ranks.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from .. import _native as N
from .norm import _finish_param_grad, _param_grad_out


def embedding_ref(ids: torch.Tensor, weight: torch.Tensor, vocab_start: int = 0) -> torch.Tensor:
    """PyTorch oracle of the (vocabulary-parallel) lookup: rows of ``weight`` for ids inside ``[vocab_start, vocab_start + V)``, zeros for the others."""
    V = weight.shape[0]
    local = ids - vocab_start
    mask = (local >= 0) & (local < V)
    out = F.embedding(local.clamp(0, V - 1), weight)
    return out * mask.unsqueeze(-1).to(out.dtype)


def embedding_backward(flat_ids: torch.Tensor, dout2d: torch.Tensor, weight: torch.Tensor, vocab_start: int = 0):
    """dW[ids - vocab_start] += dout.  Writes into ``weight.main_grad`` when present (returns None)."""
    T, H = flat_ids.numel(), weight.shape[1]
    if N.use_native(weight) and weight.dtype == torch.bfloat16 and H % 8 == 0:
        dst, acc, is_main = _param_grad_out(weight)
        if not acc:
            dst.zero_()
        sync = getattr(weight, "_stb_sparse_sync", None)
        if sync is not None and is_main and sync.fits(T, H):
            # data-parallel ranks exchange the token rows; the result already holds every rank's contribution (mean),
            # so the parameter takes no part in the dense bucket all-reduce for this gradient
            sync.accumulate(flat_ids, dout2d, dst, vocab_start, weight.shape[0])
            weight.grad_written = True
            return None
        N.call("stb_embedding_bwd", flat_ids.data_ptr(), dout2d.data_ptr(), dst.data_ptr(), T, H, vocab_start,
               weight.shape[0], N.stream_ptr())
        return _finish_param_grad(weight, dst, is_main)
    V = weight.shape[0]
    local = flat_ids - vocab_start
    mask = (local >= 0) & (local < V)
    grad = torch.zeros(weight.shape, dtype=torch.float32, device=weight.device)
    grad.index_add_(0, local.clamp(0, V - 1), dout2d.float() * mask.unsqueeze(-1))
    return grad.to(weight.dtype)


class _Embedding(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, weight, vocab_start):
        flat = ids.reshape(-1).contiguous()
        T, H = flat.numel(), weight.shape[1]
        out = torch.empty((T, H), dtype=weight.dtype, device=weight.device)
        N.call("stb_embedding_fwd", flat.data_ptr(), weight.data_ptr(), out.data_ptr(), T, H, vocab_start,
               weight.shape[0], N.stream_ptr())
        ctx.save_for_backward(flat, weight)
        ctx.vocab_start = vocab_start
        return out.view(*ids.shape, H)

    @staticmethod
    def backward(ctx, dout):
        flat, weight = ctx.saved_tensors
        T, H = flat.numel(), weight.shape[1]
        d2 = dout.reshape(T, H)
        if not d2.is_contiguous():
            d2 = d2.contiguous()
        return None, embedding_backward(flat, d2, weight, ctx.vocab_start), None


def embedding(ids: torch.Tensor, weight: torch.Tensor, vocab_start: int = 0) -> torch.Tensor:
    """Embedding lookup with an optional vocabulary shard offset: gather / scatter-add kernels for bf16 CUDA weights, the oracle otherwise."""
    if N.use_native(weight) and weight.dtype == torch.bfloat16 and weight.shape[1] % 8 == 0:
        return _Embedding.apply(ids, weight, int(vocab_start))
    return embedding_ref(ids, weight, vocab_start)
