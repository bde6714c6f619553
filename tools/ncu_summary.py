"""Markdown summary of an ncu report: one row per profiled kernel.  Usage: ncu_summary.py report.ncu-rep [more.ncu-rep]

Columns are taken from the raw page (``ncu -i rep --page raw --csv``); a metric a report lacks prints as ``-``.
Durations measured under ncu (replay, cold caches per pass) are for proportions only, never bench values.
"""
import csv
import subprocess
import sys

COLS = [
    ("time us", "gpu__time_duration.sum", 1.0),
    ("grid", "launch__grid_size", 1.0),
    ("block", "launch__block_size", 1.0),
    ("regs", "launch__registers_per_thread", 1.0),
    ("smem KB", "launch__shared_mem_per_block_dynamic", 1e-3),
    ("tensor pipe %", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", 1.0),
    ("SM busy %", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    ("DRAM %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    ("DRAM GB/s", "dram__bytes.sum.per_second", 1e-9),
    ("DRAM MB rd+wr", "dram__bytes_read.sum+dram__bytes_write.sum", 1e-6),
    ("L2 %", "lts__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    ("occupancy %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
    ("issue slots %", "sm__inst_issued.avg.pct_of_peak_sustained_active", 1.0),
    ("local ld+st", "smsp__sass_inst_executed_op_local_ld.sum+smsp__sass_inst_executed_op_local_st.sum", 1.0),
]
STALLS = "smsp__average_warps_issue_stalled_"           # ..._<reason>_per_issue_active.ratio (or warp_issue_stalled_*)


def num(s: str):
    try:
        return float(s.replace(",", ""))
    except ValueError:
        return None


def load(rep: str):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units = rows[hi], rows[hi + 1]
    return hdr, units, [r for r in rows[hi + 2:] if len(r) == len(hdr)]


def find(hdr, name):
    if name in hdr:
        return hdr.index(name)
    for i, h in enumerate(hdr):
        if h.endswith("." + name):
            return i
    return None


def scale_unit(v, unit):
    """ncu prints durations / rates in whatever unit it likes per row: normalise to us and bytes/s."""
    u = unit.strip().lower().replace("/block", "").replace("byte/s", "byte/second").replace("secondecond", "second")
    table = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6,
             "byte/second": 1.0, "kbyte/second": 1e3, "mbyte/second": 1e6, "gbyte/second": 1e9, "tbyte/second": 1e12,
             "byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}
    return v * table.get(u, 1.0)


def main() -> None:
    print("| kernel | " + " | ".join(c[0] for c in COLS) + " | top stall reasons (warps per issue) |")
    print("|---|" + "---|" * (len(COLS) + 1))
    for rep in sys.argv[1:]:
        hdr, units, rows = load(rep)
        kname = hdr.index("Kernel Name")
        stall_idx = [(i, h) for i, h in enumerate(hdr) if STALLS in h and h.endswith("per_issue_active.ratio")
                     and "_not_issued" not in h]
        for r in rows:
            cells = []
            for _, metric, mul in COLS:
                total, ok = 0.0, False
                for part in metric.split("+"):
                    i = find(hdr, part)
                    v = num(r[i]) if i is not None else None
                    if v is not None:
                        total, ok = total + scale_unit(v, units[i]), True
                cells.append(f"{total * mul:.1f}" if ok else "-")
            stalls = sorted(((num(r[i]) or 0.0, h.split(STALLS)[1].replace("_per_issue_active.ratio", "")) for i, h in stall_idx),
                            reverse=True)[:3]
            name = r[kname]
            name = name[:name.index("(")] if "(" in name else name
            print(f"| `{name[:70]}` | " + " | ".join(cells) + " | " + ", ".join(f"{n} {v:.1f}" for v, n in stalls) + " |")


if __name__ == "__main__":
    main()
