"""Top stall sites of a kernel from an ncu report (SASS view).  Usage: ncu_top_stalls.py report.ncu-rep kernel_regex [N]"""
import csv
import subprocess
import sys

rep, rx = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{rx}"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hi]
isrc, iall = hdr.index("Source"), hdr.index("Warp Stall Sampling (All Samples)")
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_")]
body = []
for r in rows[hi + 1:]:
    if len(r) <= iall or not r[0].startswith("0x"):
        if r and r[0] == "Kernel Name":
            break
        continue
    try:
        body.append((int(r[iall]), r))
    except ValueError:
        pass
total = sum(b[0] for b in body)
print(f"total samples {total}, instructions {len(body)}")
order = sorted(range(len(body)), key=lambda i: -body[i][0])[:n]
for i in sorted(order):
    s, r = body[i]
    reasons = sorted(((int(r[c]) if r[c].isdigit() else 0, hdr[c][6:]) for c in stall_cols), reverse=True)[:2]
    print(f"{i:5d} {100.0 * s / max(total, 1):5.1f}%  {r[isrc].strip()[:90]:90s} {reasons}")
