"""Per-kernel-family SASS evidence: ``python tools/dump_sass.py`` -> ``profiles/sass/<family>.sass.txt``.

Each file holds, for every instantiation of the family found in the built library, the function name, a mnemonic
census (tensor-core, TMEM, TMA, multimem, barriers, spills) and the instruction lines that carry those mnemonics with
their addresses, so ``grep -E 'UTCHMMA|LDTM|UTMALDG|UTMASTG' profiles/sass/*`` answers "does kernel X really use the
tcgen05 / TMA path" without a GPU.  The full listings are ~18 MB and are not tracked.
"""
from __future__ import annotations

import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "scaletorch_b200" / "lib" / "libscaletorch_b200.so"
OUT = ROOT / "profiles" / "sass"

FAMILIES = [
    "gemm_pair_kernel", "gemm_bf16_kernel", "grouped_gemm_kernel", "flash_attn_fwd_kernel", "attn_bwd_fused_kernel",
    "attn_bwd_dkdv_kernel", "attn_bwd_dq_kernel", "allreduce_bf16_kernel", "allreduce_f32_kernel",
    "reduce_scatter_f32_kernel", "adamw_shard_bcast_kernel", "param_bcast_kernel", "allreduce_tiny_f32_kernel",
    "ep_push_kernel", "ep_pull_kernel", "moe_router_fwd_kernel", "moe_router_bwd_kernel", "ep_plan_kernel",
    "rmsnorm_fwd_kernel", "rmsnorm_bwd_kernel", "qknorm_rope_fwd_kernel", "qknorm_rope_bwd_kernel", "swiglu",
    "stb::ce_", "adamw_kernel", "ring_", "small_allreduce", "allreduce_oneshot",
]
KEEP = re.compile(
    r"\b(UTC[A-Z]*MMA|UTCBAR|UTCCP|UTCATOM|LDTM|STTM|UTMALDG|UTMASTG|UTMAREDG|UTMAPF|UBLKCP|UBLKRED|UBLKPF|SYNCS|"
    r"MULTIMEM|LDGMC|STGMC|REDGMC|REDG|RED\b|ATOMG|MEMBAR|ERRBAR|FENCE|CCTL|UCGABAR|USETMAXREG|ACQBULK|LDG\.E\.[A-Z0-9.]*SYS|"
    r"STG\.E\.[A-Z0-9.]*SYS|LD\.E\.[A-Z0-9.]*SYS|ST\.E\.[A-Z0-9.]*SYS|LDL|STL|MUFU\.EX2)")
INSTR = re.compile(r"/\*([0-9a-f]{4,})\*/\s+(.*?);")


def main() -> int:
    if not LIB.is_file():
        print(f"{LIB} not built", file=sys.stderr)
        return 1
    txt = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    funcs: dict[str, list[str]] = {}
    name = None
    for line in txt.splitlines():
        s = line.strip()
        if s.startswith("Function :"):
            name = s.split(":", 1)[1].strip()
            funcs[name] = []
        elif name is not None:
            funcs[name].append(line)
    res: dict[str, dict[str, int]] = {}
    usage = subprocess.run(["cuobjdump", "-res-usage", str(LIB)], capture_output=True, text=True).stdout.splitlines()
    for i, line in enumerate(usage):
        m = re.match(r"\s*Function (\S+):", line)
        if m and i + 1 < len(usage):
            res[m.group(1)] = {k: int(v) for k, v in re.findall(r"(REG|STACK|SHARED|LOCAL):(\d+)", usage[i + 1])}
    demangle = subprocess.run(["c++filt"], input="\n".join(funcs), capture_output=True, text=True).stdout.splitlines()
    pretty = dict(zip(funcs, demangle))
    OUT.mkdir(parents=True, exist_ok=True)
    index = []
    for fam in FAMILIES:
        members = [f for f in funcs if fam in pretty.get(f, f)]
        if not members:
            continue
        path = OUT / f"{fam.replace('stb::', '').strip('_')}.sass.txt"
        with open(path, "w") as fh:
            fh.write(f"# cuobjdump -sass {LIB.relative_to(ROOT)} | family '{fam}' ({len(members)} instantiation(s))\n")
            fh.write("# census + the instruction lines that carry tensor-core / TMEM / TMA / multimem / barrier / spill "
                     "mnemonics (tools/dump_sass.py)\n\n")
            for f in members:
                census: collections.Counter[str] = collections.Counter()
                kept, n = [], 0
                for line in funcs[f]:
                    m = INSTR.search(line)
                    if not m:
                        continue
                    n += 1
                    k = KEEP.search(m.group(2))
                    if k:
                        op = m.group(2).split()
                        mnem = op[1] if op[0].startswith("@") and len(op) > 1 else op[0]
                        census[mnem.split(".")[0]] += 1
                        kept.append(f"  /*{m.group(1)}*/ {m.group(2)}")
                fh.write(f"Function : {pretty.get(f, f)}\n  mangled : {f}\n  instructions : {n}\n")
                fh.write("  census : " + ", ".join(f"{k}={v}" for k, v in sorted(census.items())) + "\n")
                fh.write("\n".join(kept) + "\n\n")
                index.append((fam, pretty.get(f, f)[:110], n, dict(census), res.get(f, {})))
    with open(OUT / "INDEX.md", "w") as fh:
        fh.write("# SASS census per kernel (sm_100a, `tools/dump_sass.py`)\n\n")
        fh.write("`multimem.ld_reduce` has its own opcode (`LDGMC...ADD`); `multimem.st` lowers to an ordinary "
                 "`STG.E.128.STRONG.SYS` whose address is the multicast mapping, so the NVLS broadcast kernels "
                 "(`adamw_shard_bcast_kernel<true>`, `param_bcast_kernel<true>`) show no multimem opcode here.  "
                 "`UTC*MMA` = tcgen05.mma, `LDTM` = tcgen05.ld, `UTMALDG/UTMASTG/UTMAREDG` = TMA tensor load / store / "
                 "reduce, `UBLKCP` = cp.async.bulk, `LDL+STL` = local-memory (spill) instructions.\n\n")
        fh.write("`regs` / `stack B` / `static smem B` are from `cuobjdump -res-usage` (dynamic shared memory is set at launch).\n\n")
        fh.write("| family | instantiation | instrs | regs | stack B | static smem B | UTC*MMA | LDTM | UTMALDG | UTMASTG | UTMAREDG | "
                 "UBLKCP | multimem (LDGMC/STGMC/REDGMC) | LDL+STL |\n|---|---|---|---|---|---|---|---|---|---|---|---|---|---|\n")
        for fam, nm, n, c, r in index:
            mma = sum(v for k, v in c.items() if k.startswith("UTC") and k.endswith("MMA"))
            fh.write(f"| {fam} | `{nm}` | {n} | {r.get('REG', '-')} | {r.get('STACK', '-')} | {r.get('SHARED', '-')} | "
                     f"{mma} | {c.get('LDTM', 0)} | {c.get('UTMALDG', 0)} | "
                     f"{c.get('UTMASTG', 0)} | {c.get('UTMAREDG', 0)} | {c.get('UBLKCP', 0)} | "
                     f"{c.get('LDGMC', 0) + c.get('STGMC', 0) + c.get('REDGMC', 0) + c.get('MULTIMEM', 0)} | "
                     f"{c.get('LDL', 0) + c.get('STL', 0)} |\n")
    print(f"wrote {len(index)} kernels into {OUT}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
