"""Roofline report of the fused compute+collective kernels against their unfused (GEMM + NCCL) counterparts.

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_fused.py [--tokens 8192] [--hidden 4096] [--ffn 12288]

For every op it prints (rank 0, max over ranks, CUDA events, 10 timed iterations after 3 warm-ups):
  fused ms | unfused ms | speed-up | roofline ms = max(FLOPs / measured bf16 peak, NVLink bytes per GPU / 770 GB/s measured) | fraction
Ops: all-gather→GEMM (column-parallel fwd with SP), GEMM→reduce-scatter (row-parallel fwd with SP), GEMM→all-reduce
(row-parallel fwd without SP), fp32 gradient all-reduce (DP bucket, NVLS vs peer loads vs NCCL).
JSON lines are appended to gpurun_out/fused.jsonl (copy the ones to keep into profiles/).

"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scaletorch_b200 import ops  # noqa: E402
from scaletorch_b200.runtime import fused_ops as F  # noqa: E402
from scaletorch_b200.runtime import symm  # noqa: E402
from scaletorch_b200.utils.device import get_measured_peaks  # noqa: E402

NVLINK_GBS = 770.0      # measured peer-copy bandwidth per direction on this pool (B200_PROFILING.md); nominal 900


def timeit(fn, iters: int = 10, warmup: int = 3) -> float:
    for _ in range(warmup):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def report(out, name, fused_ms, unfused_ms, flops, nvlink_bytes, peak_tflops, extra=None):
    roof = max(flops / (peak_tflops * 1e12), nvlink_bytes / (NVLINK_GBS * 1e9)) * 1e3
    rec = {"op": name, "fused_ms": round(fused_ms, 4), "unfused_ms": round(unfused_ms, 4),
           "speedup": round(unfused_ms / fused_ms, 3), "roofline_ms": round(roof, 4),
           "fraction_of_roofline": round(roof / fused_ms, 3), "world": dist.get_world_size()}
    rec.update(extra or {})
    if dist.get_rank() == 0:
        print(json.dumps(rec), flush=True)
        out.write(json.dumps(rec) + "\n")
        out.flush()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--tokens", type=int, default=8192, help="global tokens T (rows of the gathered activation)")
    ap.add_argument("--hidden", type=int, default=4096)
    ap.add_argument("--ffn", type=int, default=12288, help="full (unsharded) intermediate size")
    ap.add_argument("--grad_elems", type=int, default=1 << 26, help="fp32 elements of the gradient bucket test")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    group = dist.group.WORLD
    if not symm.group_ready(group):
        raise SystemExit("symmetric memory is not available for this group")
    peak = get_measured_peaks().get("bf16_tflops_sustained", 1368.4)
    os.makedirs("gpurun_out", exist_ok=True)
    out = open("gpurun_out/fused.jsonl", "a") if rank == 0 else None
    T, H, I = a.tokens, a.hidden, a.ffn
    Ts, Il = T // world, I // world
    torch.manual_seed(rank)

    # ---- all-gather -> GEMM: x_shard [T/w, H], W [2*I/w, H] (gate||up shard)
    xs = (torch.randn(Ts, H, device=dev) * 0.5).bfloat16()
    w1 = (torch.randn(2 * Il, H, device=dev) * 0.02).bfloat16()
    gathered = torch.empty(T, H, device=dev, dtype=torch.bfloat16)

    def unfused_ag():
        dist.all_gather_into_tensor(gathered, xs)
        return ops.gemm(gathered, w1)

    fused = timeit(lambda: F.all_gather_gemm(xs, w1, group))
    unf = timeit(unfused_ag)
    report(out, "all_gather_gemm", fused, unf, 2.0 * T * 2 * Il * H, (world - 1) * Ts * H * 2, peak,
           {"T": T, "K": H, "N": 2 * Il})

    # ---- GEMM -> reduce-scatter: y [T, I/w] x W [H, I/w] -> [T/w, H]
    y = (torch.randn(T, Il, device=dev) * 0.5).bfloat16()
    w2 = (torch.randn(H, Il, device=dev) * 0.02).bfloat16()
    rs_out = torch.empty(Ts, H, device=dev, dtype=torch.bfloat16)

    def unfused_rs():
        full = ops.gemm(y, w2)
        dist.reduce_scatter_tensor(rs_out, full)
        return rs_out

    fused = timeit(lambda: F.gemm_reduce_scatter(y, w2, group))
    unf = timeit(unfused_rs)
    report(out, "gemm_reduce_scatter", fused, unf, 2.0 * T * H * Il, (world - 1) * Ts * H * 2, peak,
           {"T": T, "K": Il, "N": H})

    # ---- GEMM -> all-reduce (no SP)
    if T * H * 2 <= symm.region_bytes(group):
        def unfused_ar():
            full = ops.gemm(y, w2)
            dist.all_reduce(full)
            return full

        fused = timeit(lambda: F.gemm_all_reduce(y, w2, group))
        unf = timeit(unfused_ar)
        report(out, "gemm_all_reduce", fused, unf, 2.0 * T * H * Il, 2.0 * (world - 1) / world * T * H * 2, peak,
               {"T": T, "K": Il, "N": H, "nvls": bool(F.get_group(group).mc_ptr)})

    # ---- fp32 gradient all-reduce (one DP bucket): NVLS vs peer loads vs NCCL
    n = a.grad_elems
    gw = F.GradWindow(group, n)
    ref = torch.randn(n, device=dev)
    nccl = timeit(lambda: dist.all_reduce(ref))
    mc = gw.mc_ptr
    for kind, ptr in (("nvls", mc), ("p2p", 0)):
        if kind == "nvls" and not mc:
            continue
        gw.mc_ptr = ptr
        for budget in (8, 16, 32, 148):
            t = timeit(lambda: gw.all_reduce_range_(0, n, 1.0 / world, budget))
            # per-GPU NVLink bytes: NVLS moves ~n*4 each way; two-shot peer access 2*(w-1)/w*n*4
            nv = n * 4.0 if kind == "nvls" else 2.0 * (world - 1) / world * n * 4.0
            report(out, f"grad_allreduce_{kind}_{budget}cta", t, nccl, 0.0, nv, peak, {"elems": n})
    gw.mc_ptr = mc
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
