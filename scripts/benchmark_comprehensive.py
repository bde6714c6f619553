#!/usr/bin/env python
"""Comprehensive parallelism benchmark on one 8×B200 node.

Counterpart of the reference's ``scripts/benchmark_comprehensive.py:54-367`` (57 labelled configurations, torchrun per
configuration, metrics parsed from logs).  Here every configuration is one ``bench.py`` run, which already prints one
machine-readable JSON line (device-timed with CUDA events, max over ranks, full optimiser step), so nothing is scraped
from console text.  Results are appended to a JSON file after every configuration (a crash loses nothing) and rendered
as a Markdown table at the end.

    python scripts/benchmark_comprehensive.py --gpus 8 --out results/bench_comprehensive.json
    python scripts/benchmark_comprehensive.py --gpus 8 --filter Qwen3-8B --impl both --steps 10 --warmup 3
    python scripts/benchmark_comprehensive.py --list
"""
from __future__ import annotations

import argparse
import json
import os
import random
import subprocess
import sys
import time
from dataclasses import asdict, dataclass

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@dataclass
class Cfg:
    label: str
    model: str
    tp: int = 1
    pp: int = 1
    cp: int = 1
    ep: int = 1
    bs: int = 1
    ga: int = 1
    seq: int = 2048
    gc: bool = False
    sp: bool = False

    def world_min(self) -> int:
        return self.tp * self.pp * self.cp * self.ep


def matrix() -> list[Cfg]:
    """The configuration families of the reference's table (dense DP / TP / SP / CP / PP mixes per model size, MoE
    EP×TP mixes), re-sized for 180 GB B200s: longer sequences, fewer forced gradient-checkpointing rows."""
    c: list[Cfg] = []
    for m, bs in (("Qwen3-0.6B", 2), ("Qwen3-1.7B", 2), ("Qwen3-4B", 1)):
        c += [Cfg(f"{m} DP", m, bs=bs), Cfg(f"{m} DP GA2", m, bs=bs, ga=2), Cfg(f"{m} DP S8192", m, bs=1, seq=8192),
              Cfg(f"{m} TP2", m, tp=2, bs=bs), Cfg(f"{m} SP-TP2", m, tp=2, sp=True, bs=bs), Cfg(f"{m} TP4", m, tp=4, bs=bs),
              Cfg(f"{m} CP2 S4096", m, cp=2, seq=4096), Cfg(f"{m} CP4 S8192", m, cp=4, seq=8192),
              Cfg(f"{m} TP2-CP2 S4096", m, tp=2, cp=2, seq=4096), Cfg(f"{m} PP2 GA4", m, pp=2, ga=4, bs=bs)]
    for m in ("Qwen3-8B", "Qwen3-14B"):
        c += [Cfg(f"{m} TP2", m, tp=2), Cfg(f"{m} SP-TP2", m, tp=2, sp=True), Cfg(f"{m} SP-TP4", m, tp=4, sp=True),
              Cfg(f"{m} SP-TP8", m, tp=8, sp=True), Cfg(f"{m} SP-TP2-CP2 S8192", m, tp=2, cp=2, sp=True, seq=8192),
              Cfg(f"{m} SP-TP4-CP2 S8192", m, tp=4, cp=2, sp=True, seq=8192),
              Cfg(f"{m} SP-TP4-PP2 GA8 S4096", m, tp=4, pp=2, sp=True, ga=8, seq=4096),
              Cfg(f"{m} SP-TP2 GC", m, tp=2, sp=True, gc=True)]
    c += [Cfg("Qwen3-32B SP-TP8", "Qwen3-32B", tp=8, sp=True), Cfg("Qwen3-32B SP-TP8 S4096", "Qwen3-32B", tp=8, sp=True, seq=4096),
          Cfg("Qwen3-32B SP-TP8 BS2 GC", "Qwen3-32B", tp=8, sp=True, bs=2, gc=True),
          Cfg("Qwen3-32B SP-TP4-PP2 GA8", "Qwen3-32B", tp=4, pp=2, sp=True, ga=8)]
    m = "Qwen3-30B-A3B"
    c += [Cfg(f"{m} EP8 S4096", m, ep=8, seq=4096), Cfg(f"{m} EP8", m, ep=8), Cfg(f"{m} EP4-TP2", m, ep=4, tp=2),
          Cfg(f"{m} EP2-TP4", m, ep=2, tp=4), Cfg(f"{m} EP2-TP4 S4096", m, ep=2, tp=4, seq=4096),
          Cfg(f"{m} SP-EP2-TP4 GC", m, ep=2, tp=4, sp=True, gc=True), Cfg(f"{m} EP8 GC BS2", m, ep=8, bs=2, gc=True)]
    return c


def command(cfg: Cfg, gpus: int, impl: str, steps: int, warmup: int, port: int) -> list[str]:
    cmd = [sys.executable]
    if gpus > 1:
        cmd += ["-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={gpus}", "--master-addr", "127.0.0.1",
                "--master-port", str(port)]
    cmd += [os.path.join(ROOT, "bench.py"), "--gpus", str(gpus), "--steps", str(steps), "--warmup", str(warmup), "--impl", impl,
            "--model", cfg.model, "--tp", str(cfg.tp), "--pp", str(cfg.pp), "--cp", str(cfg.cp), "--ep", str(cfg.ep),
            "--micro_batch_size", str(cfg.bs), "--grad_accum", str(cfg.ga), "--seq_len", str(cfg.seq)]
    if cfg.gc:
        cmd.append("--gradient_checkpointing")
    if cfg.sp:
        cmd.append("--sequence_parallel")
    return cmd


def run_one(cfg: Cfg, gpus: int, impl: str, steps: int, warmup: int, timeout: int) -> dict:
    port = 20000 + random.randint(0, 9000)
    t0 = time.time()
    rec = {"label": cfg.label, "impl": impl, "cfg": asdict(cfg), "gpus": gpus}
    try:
        p = subprocess.run(command(cfg, gpus, impl, steps, warmup, port), capture_output=True, text=True, timeout=timeout, cwd=ROOT)
        lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
        if p.returncode == 0 and lines:
            rec["result"] = json.loads(lines[-1])
            rec["status"] = "ok" if "value" in rec["result"] else "unavailable"
        else:
            rec["status"] = "oom" if "out of memory" in (p.stderr + p.stdout).lower() else "failed"
            rec["stderr_tail"] = p.stderr[-2000:]
    except subprocess.TimeoutExpired:
        rec["status"] = "timeout"
    rec["wall_s"] = round(time.time() - t0, 1)
    return rec


def markdown(records: list[dict]) -> str:
    rows = ["| config | impl | GPUs | tok/s | tok/s/GPU | ms/step | MFU % (2.25 PF) | status |", "|---|---|---|---|---|---|---|---|"]
    for r in records:
        res = r.get("result") or {}
        v = res.get("value")
        rows.append("| {} | {} | {} | {} | {} | {} | {} | {} |".format(
            r["label"], r["impl"], r["gpus"], f"{v:,.0f}" if v else "–", f"{v / r['gpus']:,.0f}" if v else "–",
            f"{res.get('ms_per_step', 0):.1f}" if v else "–",
            f"{res.get('mfu_pct_of_nominal_2.25PF', 0):.1f}" if res.get("mfu_pct_of_nominal_2.25PF") else "–", r["status"]))
    return "\n".join(rows)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "both"])
    ap.add_argument("--filter", default="", help="substring of the label")
    ap.add_argument("--timeout", type=int, default=900)
    ap.add_argument("--out", default=os.path.join(ROOT, "results", "bench_comprehensive.json"))
    ap.add_argument("--list", action="store_true")
    a = ap.parse_args()
    cfgs = [c for c in matrix() if a.filter in c.label and a.gpus % c.world_min() == 0]
    if a.list:
        for c in cfgs:
            print(c)
        print(f"{len(cfgs)} configurations")
        return
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    records: list[dict] = []
    if os.path.exists(a.out):
        records = json.load(open(a.out))
    done = {(r["label"], r["impl"]) for r in records if r["status"] == "ok"}
    for c in cfgs:
        for impl in (["ours", "reference"] if a.impl == "both" else [a.impl]):
            if (c.label, impl) in done:
                continue
            print(f"[{time.strftime('%H:%M:%S')}] {c.label} ({impl}) ...", flush=True)
            rec = run_one(c, a.gpus, impl, a.steps, a.warmup, a.timeout)
            print(f"    -> {rec['status']} {rec.get('result', {}).get('value', '')}", flush=True)
            records.append(rec)
            json.dump(records, open(a.out, "w"), indent=1)
    md = markdown(records)
    open(os.path.splitext(a.out)[0] + ".md", "w").write(md + "\n")
    print(md)


if __name__ == "__main__":
    main()
