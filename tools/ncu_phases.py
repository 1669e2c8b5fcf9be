import csv, subprocess, sys, io, collections, re
rep, want = sys.argv[1], sys.argv[2]
ranges = eval(sys.argv[3])  # list of (name, lo, hi) for msda_tile.cuh
txt = subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","cuda,sass"],capture_output=True,text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
i=0; cur=None; blocks=[]
while i < len(rows):
    r=rows[i]
    if r and r[0]=="File Path":
        cur={"file":r[1],"func":rows[i+1][1],"hdr":rows[i+2],"rows":[]}; blocks.append(cur); i+=3; continue
    if cur is not None and r: cur["rows"].append(r)
    i+=1
tot=collections.Counter(); smp=collections.Counter()
for b in blocks:
    if want not in b["func"]: continue
    h=b["hdr"]
    if "Instructions Executed" not in h: continue
    li=h.index("Line No"); ie=h.index("Instructions Executed"); ss=h.index("# Samples")
    f=b["file"].split("/")[-1]
    for r in b["rows"]:
        try: n=int(float(r[ie])); ln=int(r[li]); s=int(float(r[ss]))
        except: continue
        key=f
        if f==(sys.argv[4] if len(sys.argv)>4 else "msda_tile.cuh"):
            key="tile:other"
            for name,lo,hi in ranges:
                if lo<=ln<=hi: key="tile:"+name
        tot[key]+=n; smp[key]+=s
T=sum(tot.values()); S=sum(smp.values())
for k,v in tot.most_common(): print(f"{k:28s} {v:12d} {100*v/T:5.1f}%   samples {100*smp[k]/max(S,1):5.1f}%")
print("total",T)
