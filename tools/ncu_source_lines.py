import csv, sys, subprocess
rep, kern = sys.argv[1], sys.argv[2]
out = subprocess.run(["ncu","-i",rep,"--page","source","--print-source","cuda,sass","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(out.splitlines()))
cur=None; agg=[]; active=False
tot_s=0; tot_i=0
fname=None
for r in rows:
    if len(r)==2 and r[0]=="File Path": fname=r[1].split('/')[-1]; continue
    if len(r)==2 and r[0]=="Function Name": active = kern in r[1]; continue
    if not active: continue
    if len(r)<8: continue
    if r[0]=="Line No": continue
    if r[0]!="":   # cuda line summary row
        try:
            samples=int(r[6]); inst=int(r[7])
        except: continue
        agg.append((samples,inst,fname,r[0],r[1].strip()[:110]))
        tot_s+=samples; tot_i+=inst
agg.sort(reverse=True)
print("total samples",tot_s,"total warp inst",tot_i)
for s,i,f,l,src in agg[:int(sys.argv[3]) if len(sys.argv)>3 else 30]:
    print("%5.1f%% smp %5.1f%% inst  %s:%s  %s"%(100*s/tot_s,100*i/tot_i,f,l,src))
