"""Bisects which stage of the expert-parallel MoE layer breaks CUDA-graph capture.
    torchrun --nproc-per-node 2 --master-addr 127.0.0.1 tools/debug_capture_moe.py"""
import os
import sys
from types import SimpleNamespace

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
from scaletorch_b200 import ops  # noqa: E402
from scaletorch_b200.models.model_qwen3_moe import MoELayer  # noqa: E402
from scaletorch_b200.ops.moe_routing import route  # noqa: E402
from scaletorch_b200.parallel.expert_parallel import fused_ep  # noqa: E402

group = dist.group.WORLD
cfg = SimpleNamespace(num_experts=8, num_experts_per_tok=2, hidden_size=256, moe_intermediate_size=128, norm_topk_prob=True,
                      router_aux_loss_coef=0.01)
ctx = SimpleNamespace(sp=False, ep_size=world, ep_rank=rank, ep_group=group, tp_group=None)
torch.manual_seed(7)
layer = MoELayer(cfg, ctx).to(dev).bfloat16()
from scaletorch_b200.parallel.data_parallel.data_parallel import get_store  # noqa: E402

store = get_store(layer)          # weight gradients go to main_grad from the kernel epilogues (as in the trainer)
x = torch.randn(512, 256, device=dev).bfloat16().requires_grad_()
gout = torch.randn(512, 256, device=dev).bfloat16()


def full(bwd):
    y = layer(x)
    if bwd:
        store.zero_grad()
        return torch.autograd.grad(y, x, gout)[0]
    return y


def stage_router():
    logits = ops.linear(x.detach(), layer.router.gate.weight)
    return route(logits, 2, 0.01, True)


def stage_state():
    w, idx, pos, counts, aux = stage_router()
    T, H = x.shape
    e_local = cfg.num_experts // world
    rows_cap = (int(world * T * 2) + e_local * 127 + 127) // 128 * 128
    win = fused_ep.get_window(group, rows_cap, H, cfg.num_experts)
    return fused_ep.FusedEPState(win, idx, cfg.num_experts, (pos, counts))


def stage_push():
    st = stage_state()
    st.push(x.detach().contiguous(), None)
    return st


def stage_bwd_until(which):
    """forward, then the backward of only the first `which` autograd nodes from the output side"""
    logits = ops.linear(x, layer.router.gate.weight)
    w, idx, pos, counts, aux = route(logits, 2, 0.01, True)
    T, H = x.shape
    e_local = cfg.num_experts // world
    rows_cap = (int(world * T * 2) + e_local * 127 + 127) // 128 * 128
    win = fused_ep.get_window(group, rows_cap, H, cfg.num_experts)
    st = fused_ep.FusedEPState(win, idx, cfg.num_experts, (pos, counts))
    from scaletorch_b200.ops.activation import swiglu
    from scaletorch_b200.ops.grouped_native import _GroupedLinearPadded

    xp = fused_ep._Dispatch.apply(x, st)
    gu = _GroupedLinearPadded.apply(xp, layer.experts.gate_up_proj, st.layout, {"dx": win.y})
    act = swiglu(gu)
    y = _GroupedLinearPadded.apply(act, layer.experts.down_proj, st.layout, {"out": win.y})
    out = fused_ep._Combine.apply(y, w, st)
    store.zero_grad()
    targets = {"combine": (y, w), "down": (act,), "swiglu": (gu,), "gate_up": (xp,), "dispatch+router": (x,)}[which]
    return torch.autograd.grad(out, targets, gout, allow_unused=True)


stages = [("forward", lambda: full(False))] + [(f"bwd to {k}", (lambda k=k: stage_bwd_until(k)))
                                               for k in ("combine", "down", "swiglu", "gate_up", "dispatch+router")] + [
    ("forward+backward", lambda: full(True))]
for name, fn in stages:
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    g = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            fn()
        g.replay()
        torch.cuda.synchronize()
        res = "ok"
    except Exception as e:
        res = f"FAILED: {type(e).__name__}: {str(e).splitlines()[0][:160]}"
        torch.cuda.synchronize()
    if rank == 0:
        print(f"capture {name}: {res}", flush=True)
    del g
    dist.barrier()
dist.destroy_process_group()
