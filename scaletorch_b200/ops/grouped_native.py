"""Device-scheduled grouped GEMM path for MoE experts (``csrc/grouped_gemm.cu``).

Expert-sorted rows are scattered into a layout in which every expert's segment is padded to a multiple of 128
rows; all index math (padded offsets, tile→expert table, row destinations) is done with device tensors, so the
MoE layer runs without a single host synchronisation.  The reference synchronises twice per layer for the split
sizes and once per expert for the ``numel() == 0`` checks (``ep_comms.py:96,102``, ``model_qwen3_moe.py:143-171``).
"""
from __future__ import annotations

import ctypes

import torch

from .. import _native as N
from .activation import swiglu

_P, _I = ctypes.c_void_p, ctypes.c_int
N.register("stb_grouped_gemm_bf16", [_I, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _P])


class GroupedLayout:
    """Padded expert-major row layout derived from per-expert token counts (all on device)."""

    def __init__(self, tokens_per_expert: torch.Tensor, num_rows: int):
        E = tokens_per_expert.numel()
        dev = tokens_per_expert.device
        counts = tokens_per_expert.to(torch.int64)
        pad = (counts + 127) // 128 * 128
        zero = torch.zeros(1, dtype=torch.int64, device=dev)
        self.offs = torch.cat([zero, torch.cumsum(counts, 0)])
        offs_pad = torch.cat([zero, torch.cumsum(pad, 0)])
        self.offs_pad = offs_pad.to(torch.int32).contiguous()
        self.r_max = (num_rows + E * 127 + 127) // 128 * 128          # static upper bound: no host sync
        tiles = torch.arange(self.r_max // 128, device=dev, dtype=torch.int64) * 128
        te = torch.searchsorted(offs_pad[1:], tiles, right=True)
        self.tile_expert = torch.where(tiles < offs_pad[-1], te, torch.full_like(te, -1)).to(torch.int32).contiguous()
        rows = torch.arange(num_rows, device=dev, dtype=torch.int64)
        row_e = torch.searchsorted(self.offs[1:], rows, right=True).clamp_(max=E - 1)
        self.dest = rows - self.offs[row_e] + offs_pad[row_e]          # padded position of every sorted row
        self.num_rows, self.E = num_rows, E

    @classmethod
    def from_padded_offsets(cls, offs_pad: torch.Tensor, r_max: int, num_experts: int) -> "GroupedLayout":
        """Layout for rows that already sit at their padded positions (EP receive regions): ``offs_pad [E+1]``."""
        self = cls.__new__(cls)
        offs_pad = offs_pad.to(torch.int64)
        self.offs_pad = offs_pad.to(torch.int32).contiguous()
        self.r_max, self.E, self.num_rows = r_max, num_experts, None
        tiles = torch.arange(r_max // 128, device=offs_pad.device, dtype=torch.int64) * 128
        te = torch.searchsorted(offs_pad[1:].contiguous(), tiles, right=True)
        self.tile_expert = torch.where(tiles < offs_pad[-1], te, torch.full_like(te, -1)).to(torch.int32).contiguous()
        return self

    @classmethod
    def from_device_tables(cls, offs_pad: torch.Tensor, tile_expert: torch.Tensor, r_max: int, num_experts: int):
        """Layout whose ``offs_pad [E+1]`` / ``tile_expert [r_max / 128]`` int32 tables were written by a kernel
        (``stb_ep_plan``)."""
        self = cls.__new__(cls)
        self.offs_pad, self.tile_expert = offs_pad, tile_expert
        self.r_max, self.E, self.num_rows = r_max, num_experts, None
        return self

    def pad(self, x_sorted: torch.Tensor) -> torch.Tensor:
        out = torch.zeros((self.r_max, x_sorted.shape[1]), dtype=x_sorted.dtype, device=x_sorted.device)
        return out.index_copy(0, self.dest, x_sorted)

    def unpad(self, y_pad: torch.Tensor) -> torch.Tensor:
        return y_pad.index_select(0, self.dest)


def _call(mode, a, b, c, layout: GroupedLayout, n, k, out_fp32=0, accumulate=0):
    N.call("stb_grouped_gemm_bf16", mode, a.data_ptr(), b.data_ptr(), c.data_ptr(), layout.tile_expert.data_ptr(),
           layout.offs_pad.data_ptr(), layout.r_max, n, k, layout.E, out_fp32, accumulate, N.stream_ptr())


class _GroupedLinearPadded(torch.autograd.Function):
    """``y_pad = x_pad · W[e]ᵀ`` on padded rows; dX / dW through the dgrad / ragged-wgrad modes."""

    @staticmethod
    def forward(ctx, x_pad, weight, layout, bufs=None):
        """``bufs``: optional ``{"out": tensor, "dx": tensor}`` preallocated destinations (the EP path points them
        at its symmetric return region so the combine / dispatch-backward pull needs no staging copy).  They travel
        in a dict so autograd does not treat them as inputs."""
        E, Nn, K = weight.shape
        bufs = bufs or {}
        y = bufs.get("out")
        if y is None:
            y = torch.empty((layout.r_max, Nn), dtype=x_pad.dtype, device=x_pad.device)
        ctx.dx_out = bufs.get("dx")
        _call(0, x_pad, weight, y, layout, Nn, K)
        ctx.save_for_backward(x_pad, weight)
        ctx.layout = layout
        return y

    @staticmethod
    def backward(ctx, dy):
        x_pad, weight = ctx.saved_tensors
        layout = ctx.layout
        E, Nn, K = weight.shape
        dy = dy.contiguous()
        dx = None
        if ctx.needs_input_grad[0]:
            dx = ctx.dx_out if ctx.dx_out is not None else torch.empty_like(x_pad)
            _call(1, dy, weight, dx, layout, Nn, K)
        dw = None
        if ctx.needs_input_grad[1]:
            mg = getattr(weight, "main_grad", None)
            if mg is not None:
                acc = int(bool(getattr(weight, "grad_written", False)))
                _call(2, dy, x_pad, mg, layout, Nn, K, out_fp32=1, accumulate=acc)
                weight.grad_written = True
                hook = getattr(weight, "grad_ready_hook", None)
                if hook is not None:
                    hook()
            else:
                dw32 = torch.empty((E, Nn, K), dtype=torch.float32, device=x_pad.device)
                _call(2, dy, x_pad, dw32, layout, Nn, K, out_fp32=1, accumulate=0)
                dw = dw32.to(weight.dtype)
        return dx, dw, None, None


def supported(weight: torch.Tensor) -> bool:
    """Shape rule of the grouped GEMM: output features a multiple of 128, input features a multiple of 64."""
    _, Nn, K = weight.shape
    return Nn % 128 == 0 and K % 64 == 0


def grouped_linear_native(x_sorted, weight, tokens_per_expert):
    """Grouped linear through the padded expert-major layout (pad rows to 128 per expert, one kernel, un-pad); differentiable."""
    layout = GroupedLayout(tokens_per_expert, x_sorted.shape[0])
    return layout.unpad(_GroupedLinearPadded.apply(layout.pad(x_sorted), weight, layout))


def grouped_mlp_native(x_sorted, tokens_per_expert, gate_up_proj, down_proj):
    """SwiGLU expert MLP with ONE padded layout shared by both grouped GEMMs (pad once, un-pad once)."""
    layout = GroupedLayout(tokens_per_expert, x_sorted.shape[0])
    xp = layout.pad(x_sorted)
    gu = _GroupedLinearPadded.apply(xp, gate_up_proj, layout)
    y = _GroupedLinearPadded.apply(swiglu(gu), down_proj, layout)
    return layout.unpad(y)
