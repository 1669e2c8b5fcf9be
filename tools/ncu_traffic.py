#!/usr/bin/env python
"""Write profiles/r1_traffic.json (DRAM bytes per launch of the headline kernels) from an `ncu --set full` report of
`bench.py --no-cpu --no-extra`.  usage: tools/ncu_traffic.py report.ncu-rep out.json"""
import csv, io, json, subprocess, sys
rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
def val(r, key):
    i = hdr.index(key)
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
    return float(r[i].replace(",", "")) * mult
res = {"source": f"ncu --set full --clock-control none ({rep.split('/')[-1]}), dram__bytes_read.sum + dram__bytes_write.sum per launch"}
for r in rows[2:]:
    name = r[hdr.index("Kernel Name")]
    b = val(r, "dram__bytes_read.sum") + val(r, "dram__bytes_write.sum")
    if "fwd_lean" in name or "fwd_pipe" in name:
        res["forward_bytes_per_launch"] = b
        res["forward_kernel"] = name[:80]
    if "bwd_lean" in name:
        res["backward_kernel_bytes_per_launch"] = b
        res["backward_kernel"] = name[:80]
N, S, M, D = 64, 1875, 8, 64
memset = N * S * M * D * 4   # cudaMemsetAsync of the fp32 accumulation image: pure writes, not visible to -k regex:lean
if "backward_kernel_bytes_per_launch" in res:
    res["backward_memset_bytes"] = memset
    res["backward_bytes_per_launch"] = res["backward_kernel_bytes_per_launch"] + memset
json.dump(res, open(out, "w"), indent=1)
print(json.dumps(res, indent=1))
