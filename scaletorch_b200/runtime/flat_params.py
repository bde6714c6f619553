"""Flat parameter / gradient storage.

All trainable parameters of a model replica live in ONE contiguous buffer in the training dtype and
all gradients in ONE contiguous fp32 buffer (``main_grad``); each ``nn.Parameter`` becomes a view.
That layout is what makes the optimiser step three kernel launches, lets the wgrad GEMM epilogues
accumulate straight into ``main_grad`` (TMA reduce-add), and lets data-parallel buckets be plain
*ranges* of the gradient buffer (no pack / unpack copies).

Reference counterpart: the per-bucket flat buffers + ``param.main_grad`` views of
``scaletorch/parallel/data_parallel/bucket.py:160-356`` (there only for DP, and always fp32 by accident,
SURVEY §2.7 #2).  Here fp32 accumulation is the explicit choice (decision D4) and the layout is used
on a single GPU as well.

Segments: ``dense`` parameters (gradients reduced over DP×CP(×EP)) come first, ``expert`` parameters
(reduced over DP×CP only) second, so each reduction domain is one contiguous range.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn

ALIGN = 128  # elements; keeps every parameter 256-byte aligned in bf16 (TMA / vector loads)


@dataclass
class ParamSlot:
    """Where one parameter lives in the flat buffers: name, offset and length in elements, and its reduction segment."""
    name: str
    param: nn.Parameter
    offset: int
    numel: int
    segment: str  # "dense" | "expert"


def is_expert_param(name: str) -> bool:
    """True for parameters of MoE experts (reduced over DP x CP only, not over EP)."""
    return ".experts." in name


class FlatParamStore:
    """Re-homes every trainable parameter of a module as a view into one flat buffer (training dtype) and gives each a ``main_grad`` view into one flat fp32 gradient buffer; dense parameters first, expert parameters second."""
    def __init__(self, module: nn.Module, grad_dtype: torch.dtype = torch.float32, grad_alloc=None, param_alloc=None):
        """``grad_alloc(numel) -> fp32 tensor`` lets the caller place the gradient buffer in symmetric memory
        (data-parallel buckets are then reduced in place over NVLink); ``param_alloc(numel, dtype)`` does the same
        for the parameter buffer (the sharded optimiser writes updated parameters into every rank's copy)."""
        self.module = module
        named = []
        seen: set[int] = set()
        for name, p in module.named_parameters():
            if not p.requires_grad or id(p) in seen:
                continue
            seen.add(id(p))
            named.append((name, p))
        if not named:
            raise ValueError("module has no trainable parameters")
        self.device = named[0][1].device
        self.param_dtype = named[0][1].dtype
        for n, p in named:
            if p.dtype != self.param_dtype or p.device != self.device:
                raise ValueError(f"parameter {n} has dtype/device {p.dtype}/{p.device}; expected "
                                 f"{self.param_dtype}/{self.device} (cast the module before building the store)")
        dense = [(n, p) for n, p in named if not is_expert_param(n)]
        expert = [(n, p) for n, p in named if is_expert_param(n)]
        self.slots: list[ParamSlot] = []
        off = 0
        for seg, items in (("dense", dense), ("expert", expert)):
            for n, p in items:
                self.slots.append(ParamSlot(n, p, off, p.numel(), seg))
                off += (p.numel() + ALIGN - 1) // ALIGN * ALIGN
            if seg == "dense":
                self.dense_numel = off
        self.total_numel = off
        self.expert_numel = off - self.dense_numel
        self.param_flat = (param_alloc(off, self.param_dtype) if param_alloc is not None
                           else torch.zeros(off, dtype=self.param_dtype, device=self.device))
        assert self.param_flat.numel() >= off and self.param_flat.dtype == self.param_dtype
        self.grad_flat = grad_alloc(off) if grad_alloc is not None else torch.zeros(off, dtype=grad_dtype, device=self.device)
        assert self.grad_flat.numel() >= off and self.grad_flat.dtype == grad_dtype
        for s in self.slots:
            view = self.param_flat[s.offset:s.offset + s.numel].view(s.param.shape)
            view.copy_(s.param.data)
            s.param.data = view
            s.param.main_grad = self.grad_flat[s.offset:s.offset + s.numel]
            s.param.grad_written = False
            s.param._stb_slot = s
        self._by_param = {id(s.param): s for s in self.slots}
        # a parameter reachable under several names (tied embedding / LM head) receives one kernel-epilogue
        # gradient contribution per use; data-parallel buckets wait for all of them.  (Set here rather than in
        # the model constructor because ``to_empty`` re-creates Parameter objects and drops attributes.)
        uses: dict[int, int] = {}
        for _n, p in module.named_parameters(remove_duplicate=False):
            uses[id(p)] = uses.get(id(p), 0) + 1
        for s in self.slots:
            s.param._stb_expected_grads = uses.get(id(s.param), 1)
        self._hooks = []
        self._install_fallback_hooks()

    # Parameters whose gradient arrives through plain autograd (``.grad``) instead of a kernel epilogue
    # are folded into main_grad here, exactly like the reference's AccumulateGrad hook
    # (``data_parallel.py:254-295``): main_grad (+)= grad; grad = None.
    def _install_fallback_hooks(self) -> None:
        import os

        verbose = os.environ.get("SCALETORCH_B200_VERBOSE", "0") == "1"
        seen: set[str] = set()

        def make(p: nn.Parameter):
            def hook(param):
                if param.grad is None:
                    return
                if verbose:
                    name = getattr(getattr(param, "_stb_slot", None), "name", "?")
                    if name not in seen:
                        seen.add(name)
                        print(f"[flat_params] gradient of {name} arrived through autograd (.grad), not a kernel epilogue", flush=True)
                g = param.grad.detach().reshape(-1)
                if param.grad_written:
                    param.main_grad.add_(g.to(param.main_grad.dtype))
                else:
                    param.main_grad.copy_(g)
                    param.grad_written = True
                param.grad = None
                ready = getattr(param, "grad_ready_hook", None)
                if ready is not None:
                    ready(True)  # autograd already summed every use of the parameter

            return hook

        for s in self.slots:
            self._hooks.append(s.param.register_post_accumulate_grad_hook(make(s.param)))

    # ------------------------------------------------------------------ API
    def parameters(self) -> list[nn.Parameter]:
        return [s.param for s in self.slots]

    def slot_of(self, p: nn.Parameter) -> ParamSlot:
        return self._by_param[id(p)]

    def segment_range(self, segment: str) -> tuple[int, int]:
        return (0, self.dense_numel) if segment == "dense" else (self.dense_numel, self.total_numel)

    def zero_grad(self, set_to_zero: bool = False) -> None:
        """Marks every gradient as unwritten: the next backward *overwrites* (β=0 in the GEMM epilogue)
        instead of accumulating, so no memset is needed.  ``set_to_zero`` additionally clears the buffer."""
        for s in self.slots:
            s.param.grad_written = False
            s.param.grad = None
        if set_to_zero:
            self.grad_flat.zero_()

    def finalize_grads(self) -> None:
        """Zero the slices of parameters that received no gradient this step (unused parameters) so stale
        values never reach the optimiser."""
        for s in self.slots:
            if not s.param.grad_written:
                s.param.main_grad.zero_()

    def sync_tp_partial_grads(self) -> None:
        """Sum, over the TP group, the gradients of replicated parameters that every TP rank computes from only a
        part of the data: q_norm / k_norm weights (each rank holds a subset of the heads) and, under sequence
        parallelism, every norm weight (each rank sees a token shard).  Marked by ``mark_tp_partial_params``.  The
        reference has no such step (its SP path is incomplete)."""
        group = getattr(self, "tp_group", None)
        slots = getattr(self, "_tp_partial_slots", None)
        if slots is None:
            slots = self._tp_partial_slots = [s for s in self.slots if getattr(s.param, "_stb_tp_partial", False)]
        if group is None or not slots:
            return
        import torch.distributed as dist

        flat = torch.cat([s.param.main_grad.reshape(-1) for s in slots])
        dist.all_reduce(flat, group=group)
        off = 0
        for s in slots:
            n = s.param.main_grad.numel()
            s.param.main_grad.copy_(flat[off:off + n].view_as(s.param.main_grad))
            off += n

    def expose_grads(self) -> None:
        """Points ``param.grad`` at its ``main_grad`` view (for code that expects ``.grad``)."""
        for s in self.slots:
            s.param.grad = s.param.main_grad.view(s.param.shape)

    def remove_hooks(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks.clear()
