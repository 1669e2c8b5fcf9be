#!/usr/bin/env python
"""Static SASS instruction counts between barriers for kernels in a cubin/.so (a cheap proxy for the per-warp
instruction cost of straight-line phases).  usage: tools/sass_segments.py lib.so kernel-substring"""
import subprocess, sys, re
lib, want = sys.argv[1], sys.argv[2]
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
cur = None; segs = []; n = 0; ops = {}
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        if cur and want in cur: print(cur[:120], "total", sum(segs) + n, "segments", segs + [n])
        cur = m.group(1); segs = []; n = 0
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(.*?);", line)
    if m and cur and want in cur:
        n += 1
        if "BAR.SYNC" in m.group(1): segs.append(n); n = 0
if cur and want in cur: print(cur[:120], "total", sum(segs) + n, "segments", segs + [n])
