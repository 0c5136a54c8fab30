import csv, collections, subprocess, sys
rep=sys.argv[1]; topn=int(sys.argv[2]) if len(sys.argv)>2 else 40
out=subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","cuda,sass"],stdout=subprocess.PIPE,stderr=subprocess.DEVNULL,text=True).stdout
rows=list(csv.reader(out.splitlines()))
cur_file=None; hdr=None
agg=collections.defaultdict(float); src={}
for r in rows:
    if r and r[0]=="File Path": cur_file=r[1].split('/')[-1]
    elif r and r[0]=="Line No": hdr=r; ix=hdr.index("Instructions Executed")
    elif hdr and len(r)>ix and r[0].isdigit() and r[2]=="-":
        try: n=float(r[ix])
        except: n=0
        agg[(cur_file,int(r[0]))]+=n; src[(cur_file,int(r[0]))]=r[1]
tot=sum(agg.values())
print("total M warp-instr", round(tot/1e6,1))
byfile=collections.defaultdict(float)
for k,v in agg.items(): byfile[k[0]]+=v
print("  by file:", {f:round(100*v/tot,1) for f,v in byfile.items()})
for v,k in sorted(((v,k) for k,v in agg.items()), reverse=True)[:topn]:
    print("  %6.2f%%  %s:%d  %s"%(100*v/tot,k[0],k[1],src[k][:110]))
