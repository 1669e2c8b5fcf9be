#!/usr/bin/env python
"""Shared-memory wavefronts, instructions and stall samples per SASS opcode class from an .ncu-rep (source page).
usage: tools/ncu_sass_wavefronts.py report.ncu-rep [kernel-substring]"""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
want = sys.argv[2] if len(sys.argv) > 2 else ""
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
i = 0
while i < len(rows):
    r = rows[i]
    if r and r[0] == "Kernel Name":
        name = r[1]
        hdr = rows[i + 1]
        j = i + 2
        body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            if rows[j]:
                body.append(rows[j])
            j += 1
        i = j
        if want not in name:
            continue
        col = {h: k for k, h in enumerate(hdr)}
        agg = collections.defaultdict(lambda: [0, 0, 0, 0])
        for b in body:
            op = b[col["Source"]].split()[0] if b[col["Source"]].split() else "?"
            if op.startswith("@"):
                op = b[col["Source"]].split()[1]
            key = op.split(".")[0] + ("." + ".".join(op.split(".")[1:3]) if op.startswith(("LDS", "STS", "LDG", "STG", "RED", "ATOMS", "LDGSTS", "SHFL")) else "")
            a = agg[key]
            a[0] += int(b[col["Instructions Executed"]] or 0)
            a[1] += int(b[col["L1 Wavefronts Shared"]] or 0)
            a[2] += int(b[col["L1 Wavefronts Shared Ideal"]] or 0)
            a[3] += int(b[col["# Samples"]] or 0)
        tot = [sum(a[k] for a in agg.values()) for k in range(4)]
        print("====", name[:90], "instr %.1fM smem wavefronts %.1fM (ideal %.1fM) samples %d" % (tot[0] / 1e6, tot[1] / 1e6, tot[2] / 1e6, tot[3]))
        for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:28]:
            print("  %-22s instr %8.2fM %5.1f%%  wavefronts %8.2fM (ideal %8.2fM)  stall samples %5.1f%%" %
                  (key, a[0] / 1e6, 100 * a[0] / max(tot[0], 1), a[1] / 1e6, a[2] / 1e6, 100 * a[3] / max(tot[3], 1)))
        continue
    i += 1
