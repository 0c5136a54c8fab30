import csv, collections, subprocess, sys
rep=sys.argv[1]
out=subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","cuda,sass"],stdout=subprocess.PIPE,stderr=subprocess.DEVNULL,text=True).stdout
rows=list(csv.reader(out.splitlines()))
hdr=None; cur_file=None
names=["stall_barrier","stall_branch_resolving","stall_dispatch","stall_drain","stall_lg","stall_long_sb","stall_math","stall_membar","stall_mio","stall_misc","stall_no_inst","stall_not_selected","stall_selected","stall_short_sb","stall_sleep","stall_tex","stall_wait"]
tot=collections.Counter(); byline=collections.defaultdict(collections.Counter); src={}
for r in rows:
    if r and r[0]=="File Path": cur_file=r[1].split('/')[-1]
    elif r and r[0]=="Line No": hdr=r; idx={n:hdr.index(n) for n in names}; isamp=hdr.index("# Samples")
    elif hdr and len(r)>isamp and r[0].isdigit() and r[2]=="-":
        for n in names:
            try: v=float(r[idx[n]])
            except: v=0
            tot[n]+=v; byline[(cur_file,int(r[0]))][n]+=v
        src[(cur_file,int(r[0]))]=r[1]
s=sum(tot.values())
print("stall mix:", {k:round(100*v/s,1) for k,v in tot.most_common(10)})
top=sorted(byline.items(), key=lambda kv:-sum(kv[1].values()))[:int(sys.argv[2]) if len(sys.argv)>2 else 25]
for k,c in top:
    t=sum(c.values())
    print("%5.2f%% %s:%d %s | %s"%(100*t/s,k[0],k[1],src[k][:70],{a:round(100*b/t) for a,b in c.most_common(3)}))
