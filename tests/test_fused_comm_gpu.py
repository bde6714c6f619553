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
        elif which == "zero1_kernels":
            # reduce-scatter of a bucket into the owner's shard, tiny scalar all-reduce, sharded AdamW with parameter
            # broadcast: both transports (NVLS multimem when the fabric has it, then peer loads / stores)
            n = 1 << 20
            gw = F.GradWindow(group, n)
            pw = F.ParamWindow(group, n, torch.bfloat16)
            g = F.get_group(group)
            chunk = n // world
            lo, hi = rank * chunk, (rank + 1) * chunk
            hyper = torch.tensor([1e-2, 1 - 0.9, 1 - 0.95, 0.0], device=dev)
            for it in range(4):
                if it == 2:
                    gw.mc_ptr = 0
                    pw.mc_ptr = 0
                torch.manual_seed(1000 + 10 * it + rank)
                gw.grad.copy_(torch.randn(n, device=dev))
                ref = gw.grad.clone()
                dist.all_reduce(ref)
                ref /= world
                mine_before = gw.grad.clone()
                # two buckets that straddle shard boundaries
                gw.reduce_scatter_range_(0, n // 2 + 1024, 1.0 / world, lo, hi)
                gw.reduce_scatter_range_(n // 2 + 1024, n - (n // 2 + 1024), 1.0 / world, lo, hi)
                torch.cuda.synchronize()
                assert _rel(gw.grad[lo:hi], ref[lo:hi]) < 1e-5, _rel(gw.grad[lo:hi], ref[lo:hi])
                if lo > 0:
                    assert torch.equal(gw.grad[:lo], mine_before[:lo])          # not my shard: untouched
                pw.scratch[:3].copy_(torch.tensor([1.0 + rank, 2.0, -rank], device=dev))
                tot = pw.all_reduce_scalars_(3).clone()
                torch.cuda.synchronize()
                want = torch.tensor([world + world * (world - 1) / 2, 2.0 * world, -world * (world - 1) / 2], device=dev)
                assert torch.allclose(tot, want), (tot, want)
                # sharded AdamW: every rank starts from the same parameters; afterwards every copy holds all updated shards
                torch.manual_seed(5 + it)
                p0 = torch.randn(n, device=dev).bfloat16()
                pw.param.copy_(p0)
                m = torch.zeros(chunk, device=dev)
                v = torch.zeros(chunk, device=dev)
                from scaletorch_b200 import _native as Nt

                Nt.call("stb_adamw_shard_bcast", pw.bufs_dev, pw.mc_ptr or None, g.flag_pages_dev, lo,
                        gw.grad[lo:hi].data_ptr(), m.data_ptr(), v.data_ptr(), chunk, 0.9, 0.95, 1e-8, 0.1, None, 1.0,
                        hyper.data_ptr(), rank, world, 3, g.take_epochs(3, 2), 0, Nt.stream_ptr())
                torch.cuda.synchronize()
                gg = ref.float()
                m_ref = 0.1 * gg
                v_ref = 0.05 * gg * gg
                p_ref = p0.float() * (1 - 1e-2 * 0.1) - 1e-2 * (m_ref / 0.1) / ((v_ref / 0.05).sqrt() + 1e-8)
                assert _rel(pw.param, p_ref) < 3e-3, _rel(pw.param, p_ref)
                assert _rel(m, m_ref[lo:hi]) < 1e-5
                other = pw.param.clone()
                dist.broadcast(other, 0)
                assert torch.equal(other, pw.param.clone()), "parameter copies differ between ranks"
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

            cfg = SimpleNamespace(num_experts=max(8, 2 * world), num_experts_per_tok=2, hidden_size=256,
                                  moe_intermediate_size=128, norm_topk_prob=True, router_aux_loss_coef=0.0)
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
            # fp32 oracle: plain PyTorch MoE over ALL experts (weights gathered from every rank) on this rank's tokens;
            # expert weight gradients are the sum over the ranks' token contributions, compared on the local slice
            E, k, I = cfg.num_experts, cfg.num_experts_per_tok, cfg.moe_intermediate_size
            gu_all = [torch.empty_like(layer.experts.gate_up_proj.data) for _ in range(world)]
            dn_all = [torch.empty_like(layer.experts.down_proj.data) for _ in range(world)]
            dist.all_gather(gu_all, layer.experts.gate_up_proj.data.contiguous())
            dist.all_gather(dn_all, layer.experts.down_proj.data.contiguous())
            gu = torch.cat(gu_all).float().requires_grad_()
            dn = torch.cat(dn_all).float().requires_grad_()
            gate_w = layer.router.gate.weight.data.float().requires_grad_()
            x32 = x.detach().float().requires_grad_()
            logits = (x.detach() @ layer.router.gate.weight.data.t()).float()   # routing decisions as the layer makes them
            tv, ti = torch.topk(logits, k, dim=-1)
            logits32 = x32 @ gate_w.t()
            wts = torch.softmax(logits32.gather(1, ti), dim=-1)
            wts = wts / wts.sum(-1, keepdim=True)
            y32 = torch.zeros_like(x32)
            for e in range(E):
                tok, slot = (ti == e).nonzero(as_tuple=True)
                if tok.numel() == 0:
                    continue
                h = x32[tok] @ gu[e].t()
                a = torch.nn.functional.silu(h[:, :I]) * h[:, I:]
                y32 = y32.index_add(0, tok, (a @ dn[e].t()) * wts[tok, slot].unsqueeze(-1))
            y32.backward(gout.float())
            assert _rel(yf, y32) < 2e-2, _rel(yf, y32)
            assert _rel(xf, x32.grad) < 3e-2, _rel(xf, x32.grad)
            g_gu, g_dn = gu.grad.clone(), dn.grad.clone()
            dist.all_reduce(g_gu)
            dist.all_reduce(g_dn)
            el = E // world
            assert _rel(gf["experts.gate_up_proj"], g_gu[rank * el:(rank + 1) * el]) < 3e-2
            assert _rel(gf["experts.down_proj"], g_dn[rank * el:(rank + 1) * el]) < 3e-2
            assert _rel(gf["router.gate.weight"], gate_w.grad) < 5e-2, _rel(gf["router.gate.weight"], gate_w.grad)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("which", ["barrier_allreduce", "gemm_rs", "ag_gemm", "ep_moe", "zero1_kernels"])
def test_fused_collectives(which, world):
    """Every fused compute+collective kernel at 2, 4 and 8 ranks (whatever the box has)."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    mp.spawn(_worker, args=(world, _free_port(), which), nprocs=world, join=True)
