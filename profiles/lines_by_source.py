"""Executed warp instructions per CUDA source line of the WalkText kernels in an ncu report
(captured with --import-source on; the library is built with -lineinfo):
    python profiles/lines_by_source.py gpurun_out/<name>.ncu-rep [top N]"""
import csv, collections, subprocess, sys
rep=sys.argv[1]; topn=int(sys.argv[2]) if len(sys.argv)>2 else 30
out=subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","cuda,sass"],stdout=subprocess.PIPE,stderr=subprocess.DEVNULL,text=True).stdout
rows=list(csv.reader(out.splitlines()))
cur_file=None; cur_fn=None; hdr=None
agg=collections.defaultdict(float); src={}
for r in rows:
    if r and r[0]=="File Path": cur_file=r[1].split('/')[-1]
    elif r and r[0]=="Function Name": cur_fn=r[1]
    elif r and r[0]=="Line No": hdr=r; ix=hdr.index("Instructions Executed")
    elif hdr and len(r)>ix and r[0].isdigit() and r[2]=="-":
        try: n=float(r[ix])
        except: n=0
        key=(("emit" if "(bool)1" in cur_fn else "count"), cur_file, int(r[0]))
        agg[key]+=n; src[(cur_file,int(r[0]))]=r[1]
for which in ("count","emit"):
    tot=sum(v for k,v in agg.items() if k[0]==which)
    print(which, "total M warp-instr (both launches)", round(tot/1e6,1))
    byfile=collections.defaultdict(float)
    for k,v in agg.items():
        if k[0]==which: byfile[k[1]]+=v
    print("  by file:", {f:round(100*v/tot,1) for f,v in byfile.items()})
    top=sorted(((v,k) for k,v in agg.items() if k[0]==which), reverse=True)[:topn]
    for v,k in top:
        print("  %6.2f%%  %s:%d  %s"%(100*v/tot,k[1],k[2],src[(k[1],k[2])][:100]))
