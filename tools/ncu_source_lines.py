"""Developer tool: per-source-line executed warp instructions of one kernel from an .ncu-rep
(captured with --import-source on).  usage: python tools/ncu_source_lines.py rep.ncu-rep [kernel-substr] [top]"""
import csv, subprocess, sys
rep = sys.argv[1]; kern = sys.argv[2] if len(sys.argv) > 2 else ""; top = int(sys.argv[3]) if len(sys.argv) > 3 else 60
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
i = 0; sections = []
while i < len(rows):
    if rows[i] and rows[i][0] == "Function Name":
        name = rows[i][1]; hdr = rows[i + 1]; j = i + 2; body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "File Path"):
            body.append(rows[j]); j += 1
        sections.append((name, hdr, body)); i = j
    else:
        i += 1
for name, hdr, body in sections:
    if kern not in name: continue
    ci = hdr.index("Instructions Executed"); cs = hdr.index("# Samples")
    lines = [(int(r[0]), r[1], int(r[ci]) if r[ci].isdigit() else 0, int(r[cs]) if r[cs].isdigit() else 0) for r in body if r and r[0].isdigit()]
    tot = sum(l[2] for l in lines); tots = sum(l[3] for l in lines)
    print(f"== {name}: {tot} warp instructions, {tots} samples")
    for ln, src, n, smp in sorted(lines, key=lambda l: -l[2])[:top]:
        print(f"{ln:5d} {n:11d} {100*n/max(tot,1):5.1f}%  smp {100*smp/max(tots,1):5.1f}%  {src.strip()[:110]}")
