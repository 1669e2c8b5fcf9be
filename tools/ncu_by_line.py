#!/usr/bin/env python
"""Executed warp instructions (and shared-memory wavefronts) per SOURCE LINE of one kernel: joins the SASS view of an
.ncu-rep with the line table of the matching cubin (nvdisasm -g), instruction by instruction.
usage: tools/ncu_by_line.py report.ncu-rep object-or-cubin mangled-kernel-substring [top-n]"""
import csv, io, subprocess, sys, collections, re, os, tempfile
rep, obj, want = sys.argv[1], sys.argv[2], sys.argv[3]
topn = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
if not obj.endswith(".cubin"):
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
    obj = os.path.join(tmp, [f for f in os.listdir(tmp) if f.endswith(".cubin")][0])
sass = subprocess.run(["nvdisasm", "-g", "-c", obj], capture_output=True, text=True).stdout.splitlines()
# line table of the kernel: list of (file, line) per instruction, in address order
lines, on, cur = [], False, ("?", 0)
for ln in sass:
    if ln.startswith(".text."):
        on = want in ln
        continue
    if not on:
        continue
    m = re.search(r'//## File "(.*?)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
        lines.append(cur)
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
demangled = None
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name, hdr = rows[i][1], rows[i + 1]
        j = i + 2
        body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            if rows[j]:
                body.append(rows[j])
            j += 1
        i = j
        if len(body) != len(lines):
            continue
        col = {h: k for k, h in enumerate(hdr)}
        agg = collections.defaultdict(lambda: [0, 0, 0])
        for b, key in zip(body, lines):
            a = agg[key]
            a[0] += int(b[col["Instructions Executed"]] or 0)
            a[1] += int(b[col["L1 Wavefronts Shared"]] or 0)
            a[2] += int(b[col["# Samples"]] or 0)
        tot = [sum(a[k] for a in agg.values()) for k in range(3)]
        print("====", name[:100], "instr %.1fM wavefronts %.1fM" % (tot[0] / 1e6, tot[1] / 1e6))
        for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:topn]:
            print("  %-22s:%-5d instr %8.2fM %5.1f%%  smem wavefronts %7.2fM  stall samples %5.1f%%" %
                  (key[0], key[1], a[0] / 1e6, 100 * a[0] / tot[0], a[1] / 1e6, 100 * a[2] / max(tot[2], 1)))
        break
    i += 1
else:
    print("no kernel of the report has", len(lines), "instructions")
