import csv, sys, subprocess
rep, kern = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu","-i",rep,"--page","source","--print-source","cuda,sass","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=None; fname=None; agg={}
for r in rows:
    if len(r)==2 and r[0]=="File Path": fname=r[1].split('/')[-1]; continue
    if len(r)>8 and r[0]=="Line No": hdr=r; continue
    if len(r)<8 or r[0]=="": continue
    try: ln=int(r[0]); samples=int(r[6]); inst=int(r[7])
    except: continue
    agg[(fname,ln)]=(samples,inst,r[1].strip()[:90])
tot=sum(v[0] for v in agg.values()); toti=sum(v[1] for v in agg.values())
def rng(f,a,b): 
    s=sum(v[0] for (ff,l),v in agg.items() if ff==f and a<=l<=b); i=sum(v[1] for (ff,l),v in agg.items() if ff==f and a<=l<=b)
    return 100*s/tot,100*i/toti
import json
for name,f,a,b in json.loads(sys.argv[3]):
    print("%-40s samples %5.1f%%  inst %5.1f%%"%((name,)+rng(f,a,b)))
others={}
for (f,l),v in agg.items():
    others[f]=others.get(f,0)+v[0]
print({k:round(100*v/tot,1) for k,v in others.items()})
