"""Fused compute+collective kernels over NVLink peer memory vs their NCCL + GEMM equivalents (needs >= 2 GPUs;
run with `gpurun --gpus 2 -- python -m pytest tests/test_fused_comm_gpu.py -m gpu`)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def _worker(rank, world, port, which):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from scaletorch_b200 import ops
        from scaletorch_b200.runtime import fused_ops as F

        group = dist.group.WORLD
        os.environ["SCALETORCH_B200_POISON"] = "1"      # unwritten output rows would surface as NaN
        torch.manual_seed(100 + rank)
        dev = torch.device("cuda", rank)
        if which == "barrier_allreduce":
            g = F.get_group(group, 64 << 20)
            for _ in range(3):
                g.barrier()
            for n in (8 * 1024, 1 << 20):
                t = torch.randn(n, device=dev).bfloat16()
                ref = t.clone()
                dist.all_reduce(ref)
                out = F.all_reduce_(t, group)
                assert _rel(out, ref) < 1e-2, _rel(out, ref)
            # both transports of the bf16 all-reduce (NVLS multimem when the fabric has it, one-shot peer loads)
            for no_mm in ("0", "1"):
                os.environ["SCALETORCH_B200_NO_MULTIMEM"] = no_mm
                t = torch.randn(1 << 18, device=dev).bfloat16()
                ref = t.clone()
                dist.all_reduce(ref)
                assert _rel(F.all_reduce_(t, group), ref) < 1e-2
            os.environ["SCALETORCH_B200_NO_MULTIMEM"] = "0"
            gw = F.GradWindow(group, 1 << 20)
            if rank == 0:
                print(f"multicast available: {bool(g.mc_ptr)} (window) {bool(gw.mc_ptr)} (grad window)", flush=True)
            for it in range(4):
                if it == 2:
                    gw.mc_ptr = 0            # second half: two-shot peer-memory path
                gw.grad.copy_(torch.randn(1 << 20, device=dev))
                ref = gw.grad.clone()
                dist.all_reduce(ref)
                ref /= world
                gw.all_reduce_range_(0, 1 << 19, 1.0 / world)
                gw.all_reduce_range_(1 << 19, 1 << 19, 1.0 / world)
                torch.cuda.synchronize()
                assert _rel(gw.grad, ref) < 1e-5, _rel(gw.grad, ref)
        elif which == "gemm_rs":
            T, K, Nn = 512 * world, 1024, 768
            x = (torch.randn(T, K, device=dev) * 0.5).bfloat16()
            w = (torch.randn(Nn, K, device=dev) * 0.05).bfloat16()
            for _ in range(2):
                y = F.gemm_reduce_scatter(x, w, group)
                full = x.float() @ w.float().t()
                dist.all_reduce(full)
                ref = full[rank * (T // world):(rank + 1) * (T // world)]
                assert _rel(y, ref) < 2e-2, _rel(y, ref)
            # dgrad layout (B MN-major)
            wt = w.t().contiguous()     # [K, N]
            y = F.gemm_reduce_scatter(x, wt, group, b_mn=True)
            assert _rel(y, ref) < 2e-2
            y2 = F.gemm_all_reduce(x, w, group)
            assert _rel(y2, full) < 2e-2
        elif which == "ag_gemm":
            Ts, K, Nn = 512, 1024, 1536
            xs = (torch.randn(Ts, K, device=dev) * 0.5).bfloat16()
            w = (torch.randn(Nn, K, device=dev) * 0.05).bfloat16()
            gathered = torch.empty(world * Ts, K, device=dev, dtype=torch.bfloat16)
            dist.all_gather_into_tensor(gathered, xs)
            ref = gathered.float() @ w.float().t()
            for _ in range(3):
                y, x_full = F.all_gather_gemm(xs, w, group)
                assert torch.equal(x_full, gathered)
                assert _rel(y, ref) < 2e-2, _rel(y, ref)
        elif which == "ep_moe":
            from types import SimpleNamespace

            from scaletorch_b200.models.model_qwen3_moe import MoELayer

            cfg = SimpleNamespace(num_experts=8, num_experts_per_tok=2, hidden_size=256, moe_intermediate_size=128,
                                  norm_topk_prob=True, router_aux_loss_coef=0.0)
            ctx = SimpleNamespace(sp=False, ep_size=world, ep_rank=rank, ep_group=group, tp_group=None)
            torch.manual_seed(7)
            layer = MoELayer(cfg, ctx).to(dev).bfloat16()
            torch.manual_seed(50 + rank)
            x = (torch.randn(600, 256, device=dev)).bfloat16().requires_grad_()
            gout = torch.randn(600, 256, device=dev).bfloat16()
            res = {}
            for fused in (True, False, True):
                layer.use_fused_ep = fused
                layer.zero_grad()
                x.grad = None
                y = layer(x)
                y.backward(gout)
                res[fused] = (y.detach().clone(), x.grad.clone(), {n: p.grad.clone() for n, p in layer.named_parameters()})
            assert layer.dropped_pairs is not None and int(layer.dropped_pairs) == 0
            yf, xf, gf = res[True]
            yr, xr, gr = res[False]
            assert _rel(yf, yr) < 2e-2, _rel(yf, yr)
            assert _rel(xf, xr) < 3e-2, _rel(xf, xr)
            for n in gf:
                assert _rel(gf[n], gr[n]) < 3e-2, (n, _rel(gf[n], gr[n]))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("which", ["barrier_allreduce", "gemm_rs", "ag_gemm", "ep_moe"])
def test_fused_collectives_two_gpus(which):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), which), nprocs=world, join=True)
