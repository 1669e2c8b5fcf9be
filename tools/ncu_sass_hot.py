#!/usr/bin/env python
"""Hottest SASS instructions (by warp-stall samples) of a kernel in an .ncu-rep, with the dominant stall reason.
usage: tools/ncu_sass_hot.py report.ncu-rep [top-n]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1]; hdr = rows[i + 1]; i += 2
        body = []
        while i < len(rows) and not (rows[i] and rows[i][0] == "Kernel Name"):
            if rows[i]: body.append(rows[i])
            i += 1
        si, ss, ie = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
        stall_cols = [(k, h) for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        tot = sum(int(r[ss]) for r in body)
        print("====", name[:110], "| total samples", tot, "| sass instrs", len(body))
        idx = sorted(range(len(body)), key=lambda k: -int(body[k][ss]))[:topn]
        for k in sorted(idx):
            r = body[k]
            st = sorted(((int(r[c]), h[6:]) for c, h in stall_cols), reverse=True)[:2]
            print(f"{k:5d} {int(r[ss]):6d} {100.0*int(r[ss])/max(tot,1):5.1f}%  ex {int(r[ie]):9d}  {r[si].strip()[:70]:70s} {st[0][1]}:{st[0][0]} {st[1][1]}:{st[1][0]}")
    else:
        i += 1
