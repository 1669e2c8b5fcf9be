#!/usr/bin/env python
"""Per-source-line instruction counts / stall samples from an .ncu-rep (needs -lineinfo and --import-source on).
usage: tools/ncu_lines.py report.ncu-rep [kernel-substring] [top-n]"""
import csv, subprocess, sys, io, collections
rep = sys.argv[1]
want = sys.argv[2] if len(sys.argv) > 2 else ""
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
# the dump is a sequence of blocks: "File Path", "Function Name", header row, data rows (cuda view then sass view)
i = 0
blocks = []
cur = None
while i < len(rows):
    r = rows[i]
    if r and r[0] == "File Path":
        cur = {"file": r[1], "func": rows[i + 1][1], "hdr": rows[i + 2], "rows": []}
        blocks.append(cur)
        i += 3
        continue
    if cur is not None and r:
        cur["rows"].append(r)
    i += 1
for b in blocks:
    if want not in b["func"]:
        continue
    h = b["hdr"]
    if "Line No" not in h or "Instructions Executed" not in h:
        continue
    li, si = h.index("Line No"), h.index("Source")
    ie, ss = h.index("Instructions Executed"), h.index("# Samples")
    tot = 0
    agg = []
    for r in b["rows"]:
        try:
            n = int(float(r[ie])); smp = int(float(r[ss]))
        except (ValueError, IndexError):
            continue
        tot += n
        agg.append((n, smp, r[li], r[si].strip()[:110]))
    if not agg:
        continue
    print("====", b["func"][:90], "|", b["file"].split("/")[-1], "| total instr", tot)
    agg.sort(reverse=True)
    for n, smp, ln, src in agg[:topn]:
        print(f"{n:12d} {100.0 * n / max(tot, 1):5.1f}%  smp {smp:6d}  L{ln:>4s}  {src}")
