"""Stall samples per warp ROLE of a warp-specialised kernel, from the SASS source page of an ncu report:
  ncu -i rep.ncu-rep --page source --csv --print-source sass > src.csv ; python profiles/role_stalls.py src.csv [region [top N]]
Regions are split at the USETMAXREG instructions; the hot loop of a region = the instructions executed about once per
32-point step (50 536 steps per SM sub-partition sum at cfg3: adjust `mx` for another workload)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
I = {h:i for i,h in enumerate(hdr)}
data = rows[2:]
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
recs=[]
for r in data:
    if len(r) < len(hdr): continue
    try:
        addr=int(r[I["Address"]],16) if r[I["Address"]].startswith("0x") else int(r[I["Address"]])
    except: continue
    n=int(r[I["# Samples"]] or 0); ex=int(r[I["Instructions Executed"]] or 0)
    st={s[6:]:int(r[I[s]] or 0) for s in stalls}
    recs.append((addr,r[I["Source"]],n,ex,st))
base=recs[0][0]
# find USETMAXREG positions
marks=[i for i,x in enumerate(recs) if "USETMAXREG" in x[1]]
print("USETMAXREG at idx", marks, "total instrs", len(recs), "total samples", sum(x[2] for x in recs))
# identify hot loop bodies: instructions with exec count == max-ish
import collections
def summarize(lo,hi,name):
    seg=recs[lo:hi]
    tot=sum(x[2] for x in seg)
    mx=50536
    hot=[x for x in seg if 0.5*mx<=x[3]<=1.3*mx]
    spin=[x for x in seg if x[3]>1.3*mx]
    print('   spin-loop instrs',len(spin),'samples',sum(x[2] for x in spin))
    agg=collections.Counter()
    for x in hot:
        for k,v in x[4].items(): agg[k]+=v
    hs=sum(x[2] for x in hot)
    print(f"{name}: instrs {len(seg)} samples {tot}; hot loop: {len(hot)} instrs exec~{mx} samples {hs}: "+", ".join(f"{k} {100*v/max(hs,1):.0f}%" for k,v in agg.most_common(7)))
    return hot
bounds=[0]+marks+[len(recs)]
hots=[]
for i in range(len(bounds)-1):
    hots.append(summarize(bounds[i],bounds[i+1],f"region{i}"))
if len(sys.argv)>2:
    reg=int(sys.argv[2]); top=int(sys.argv[3]) if len(sys.argv)>3 else 40
    hot=hots[reg]
    for x in sorted(hot,key=lambda x:-x[2])[:top]:
        why=sorted(x[4].items(),key=lambda kv:-kv[1])[:3]
        print(f"  {x[0]-base:#06x} {x[2]:4d} {x[1][:70]:70s} "+", ".join(f"{k} {v}" for k,v in why if v))
