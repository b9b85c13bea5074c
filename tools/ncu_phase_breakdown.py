"""Developer tool (CPU): warp instructions, active lanes and stall samples of one kernel by PHASE, from the SASS-level
source page of an ncu report (`--set full --import-source on`).  Phases are address ranges of that particular build
(relative to the kernel's first instruction), found by looking at the SASS: name=0xSTART ... in ascending order.

    python tools/ncu_phase_breakdown.py rep.ncu-rep k1_search <tasks> setup=0x0 scan_head=0x12f0 ... > profiles/..._phases.txt
"""
import csv
import io
import subprocess
import sys

rep, kern, ntask = sys.argv[1], sys.argv[2], sys.argv[3]  # <tasks>: a number, or @0xADDR = executions of that instruction
marks = [(a.split("=")[0], int(a.split("=")[1], 16)) for a in sys.argv[4:]]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "-k", "regex:" + kern],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
h = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr, body = rows[h], rows[h + 1:]
ia, ie, it, ism = (hdr.index(k) for k in ("Address", "Instructions Executed", "Thread Instructions Executed", "# Samples"))
base = int(body[0][ia], 16)
data = [(int(r[ia], 16) - base, int(r[ie] or 0), int(r[it] or 0), int(r[ism] or 0)) for r in body if len(r) > it and r[ia].startswith("0x")]
ntask = float(next(d[1] for d in data if d[0] == int(ntask[1:], 16))) if ntask.startswith("@") else float(ntask)
tot, tots = sum(d[1] for d in data), sum(d[3] for d in data)
print(f"{kern}: {tot} warp instructions, {sum(d[2] for d in data) / tot:.2f} active lanes per instruction, {tots} stall samples, {ntask:.0f} tasks")
print(f"{'phase':36s} {'M inst':>8s} {'share':>6s} {'lanes':>6s} {'per task':>9s} {'samples':>8s}")
for i, (name, a0) in enumerate(marks):
    a1 = marks[i + 1][1] if i + 1 < len(marks) else 1 << 40
    e = sum(d[1] for d in data if a0 <= d[0] < a1)
    t = sum(d[2] for d in data if a0 <= d[0] < a1)
    sm = sum(d[3] for d in data if a0 <= d[0] < a1)
    print(f"{name:36s} {e / 1e6:8.2f} {100 * e / tot:5.1f}% {t / max(e, 1):6.1f} {e / ntask:9.1f} {100 * sm / max(tots, 1):7.1f}%")
