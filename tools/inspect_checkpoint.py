#!/usr/bin/env python
"""What is in a training checkpoint directory?

    python tools/inspect_checkpoint.py work_dir/300 [--json] [--tensors]

Reads every ``weights_*.pth`` shard of ``work_dir/<step>/`` (``utils/checkpoint.py::CheckpointManager``) on the CPU and
prints: step, trained tokens, the layout the files were written under and whether the set of shard files is complete for
it, parameter count and bytes per shard, optimiser state size, non-finite values in the weights, the data-loader position
and whether a scheduler state is present.  ``--tensors`` lists every tensor with shape and dtype.  Exit code 1 when the
directory is incomplete or holds NaN / Inf weights -- usable as a gate before a long resume.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import sys

import torch

NAME = re.compile(r"weights_tp_rank_world_size=(\d+)_(\d+)_pp_rank_world_size=(\d+)_(\d+)(?:_ep_rank_world_size=(\d+)_(\d+))?\.pth$")


def _tensors(obj, prefix=""):
    if torch.is_tensor(obj):
        yield prefix, obj
    elif isinstance(obj, dict):
        for k, v in obj.items():
            yield from _tensors(v, f"{prefix}.{k}" if prefix else str(k))
    elif isinstance(obj, (list, tuple)):
        for i, v in enumerate(obj):
            yield from _tensors(v, f"{prefix}[{i}]")


def inspect(path: str, list_tensors: bool = False) -> dict:
    files = sorted(f for f in os.listdir(path) if NAME.match(f))
    if not files:
        raise SystemExit(f"{path}: no weights_*.pth shard files")
    shards, seen, layout = [], set(), None
    for f in files:
        m = NAME.match(f)
        tp_r, tp, pp_r, pp, ep_r, ep = (int(x) if x is not None else None for x in m.groups())
        ep_r, ep = ep_r or 0, ep or 1
        layout = layout or {"tp": tp, "pp": pp, "ep": ep}
        seen.add((tp_r, pp_r, ep_r))
        ck = torch.load(os.path.join(path, f), map_location="cpu", weights_only=False)
        params = list(_tensors(ck.get("model", {})))
        opt = list(_tensors(ck.get("optimizer") or {}))
        bad = [n for n, t in params if t.is_floating_point() and not torch.isfinite(t).all()]
        entry = {"file": f, "tp_rank": tp_r, "pp_rank": pp_r, "ep_rank": ep_r,
                 "trained_steps": ck.get("trained_steps"), "trained_tokens": ck.get("trained_tokens"),
                 "parameters": sum(t.numel() for _, t in params), "parameter_bytes": sum(t.numel() * t.element_size() for _, t in params),
                 "optimizer_bytes": sum(t.numel() * t.element_size() for _, t in opt), "nonfinite_tensors": bad,
                 "layout_recorded": ck.get("layout"), "data_loader": (ck.get("extra") or {}).get("data_loader"),
                 "has_rng": bool(ck.get("rng")), "moments_can_be_recut": bool(ck.get("optimizer_slots"))}
        if list_tensors:
            entry["tensors"] = [{"name": n, "shape": list(t.shape), "dtype": str(t.dtype).replace("torch.", "")} for n, t in params]
        shards.append(entry)
    want = {(t, p, e) for t in range(layout["tp"]) for p in range(layout["pp"]) for e in range(layout["ep"])}
    steps = {s["trained_steps"] for s in shards}
    return {"path": path, "layout": layout, "shards_found": len(files), "shards_expected": len(want),
            "missing_shards": sorted(want - seen), "temporary_files": sorted(f for f in os.listdir(path) if f.endswith(".tmp")),
            "consistent_step": len(steps) == 1, "trained_steps": shards[0]["trained_steps"], "trained_tokens": shards[0]["trained_tokens"],
            "scheduler_state": os.path.isfile(os.path.join(path, "scheduler.pt")),
            "total_parameters_over_shards": sum(s["parameters"] for s in shards),
            "nonfinite": any(s["nonfinite_tensors"] for s in shards), "shards": shards}


def healthy(info: dict) -> bool:
    return (not info["missing_shards"] and not info["temporary_files"] and info["consistent_step"] and not info["nonfinite"])


def render(info: dict) -> str:
    lay = info["layout"]
    out = [f"{info['path']}: step {info['trained_steps']}, {info['trained_tokens']} tokens, written under tp{lay['tp']} pp{lay['pp']} ep{lay['ep']}",
           f"  shards {info['shards_found']} / {info['shards_expected']}" + (f"  MISSING {info['missing_shards']}" if info["missing_shards"] else "")
           + (f"  TEMPORARY FILES {info['temporary_files']}" if info["temporary_files"] else "")
           + ("" if info["consistent_step"] else "  SHARDS DISAGREE ON THE STEP"),
           f"  scheduler state: {'yes' if info['scheduler_state'] else 'no'}   data position: {info['shards'][0]['data_loader']}",
           "  resume under another layout (--resume_reshard): weights"
           + (" and AdamW moments" if all(s["moments_can_be_recut"] for s in info["shards"]) else " only (no optimiser slot table in these files)")]
    for s in info["shards"]:
        out.append(f"  tp{s['tp_rank']} pp{s['pp_rank']} ep{s['ep_rank']}: {s['parameters'] / 1e6:9.2f} M parameters, {s['parameter_bytes'] / 1e6:9.1f} MB weights, "
                   f"{s['optimizer_bytes'] / 1e6:9.1f} MB optimiser state" + (f"  NON-FINITE: {s['nonfinite_tensors'][:3]}" if s["nonfinite_tensors"] else ""))
        for t in s.get("tensors", []):
            out.append(f"      {t['name']:60s} {str(t['shape']):24s} {t['dtype']}")
    out.append("OK: complete and finite" if healthy(info) else "NOT USABLE AS IT IS")
    return "\n".join(out)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("path", help="checkpoint directory (work_dir/<step>)")
    ap.add_argument("--json", action="store_true")
    ap.add_argument("--tensors", action="store_true", help="list every parameter tensor")
    a = ap.parse_args(argv)
    info = inspect(a.path, a.tensors)
    print(json.dumps(info, default=str) if a.json else render(info))
    return 0 if healthy(info) else 1


if __name__ == "__main__":
    sys.exit(main())
