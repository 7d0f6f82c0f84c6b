"""Warp-stall reasons per CUDA source line from an ncu source page:
    ncu -i rep.ncu-rep --page source --csv --print-source cuda,sass --launch-skip K --launch-count 1 > src.csv
    python profiles/ncu_stalls.py src.csv [lines per reason]"""
import csv, sys, collections
rows=list(csv.reader(open(sys.argv[1])))
hdr=None; cur=None
tot=collections.Counter(); by=collections.defaultdict(collections.Counter); src={}
for r in rows:
    if len(r)==2 and r[0]=="File Path": cur=r[1].split('/')[-1]; continue
    if len(r)>40 and r[0]=="Line No": hdr=r; continue
    if hdr and len(r)==len(hdr) and r[0].isdigit() and r[2]=='-':   # CUDA source line rows
        key=(cur,int(r[0])); src[key]=r[1].strip()[:90]
        for i,h in enumerate(hdr):
            if h.startswith("stall_") and "Not Issued" not in h:
                try: n=int(r[i])
                except: continue
                tot[h]+=n; by[h][key]+=n
T=sum(tot.values())
print("total samples",T)
for h,n in tot.most_common(10):
    print("%-24s %5.1f%%"%(h,100*n/T))
    for k,m in by[h].most_common(int(sys.argv[2]) if len(sys.argv)>2 else 6):
        print("      %5.1f%%  %s:%d  %s"%(100*m/T,k[0],k[1],src[k]))
