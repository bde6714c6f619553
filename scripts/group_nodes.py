#!/usr/bin/env python
"""Group ``ip<TAB>rack`` lines by rack and print one node list per rack, largest rack first, so a job can be placed
inside as few racks (NVLink / leaf-switch domains) as possible (reference: ``scripts/group_nodes.py``).

    python scripts/group_nodes.py nodes.tsv [--need N]
"""
from __future__ import annotations

import argparse
from collections import defaultdict


def group(lines):
    racks: dict[str, list[str]] = defaultdict(list)
    for ln in lines:
        ln = ln.strip()
        if not ln or ln.startswith("#"):
            continue
        parts = ln.split()
        racks[parts[1] if len(parts) > 1 else "unknown"].append(parts[0])
    return sorted(racks.items(), key=lambda kv: (-len(kv[1]), kv[0]))


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("file")
    ap.add_argument("--need", type=int, default=0, help="print a placement of this many nodes using the fewest racks")
    a = ap.parse_args()
    racks = group(open(a.file))
    for rack, ips in racks:
        print(f"# rack {rack}: {len(ips)} nodes")
        for ip in ips:
            print(ip)
    if a.need:
        chosen: list[str] = []
        for _, ips in racks:
            chosen += ips[: a.need - len(chosen)]
            if len(chosen) >= a.need:
                break
        print(f"# placement for {a.need} nodes")
        print("\n".join(chosen))


if __name__ == "__main__":
    main()
