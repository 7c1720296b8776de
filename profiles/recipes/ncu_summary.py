#!/usr/bin/env python
"""Summarise one ncu report: key metrics + dynamic SASS opcode mix per ray.  usage: ncu_summary.py rep nrays [launch index]"""
import csv, collections, subprocess, sys, io
rep, nrays = sys.argv[1], int(sys.argv[2])
launch = int(sys.argv[3]) if len(sys.argv) > 3 else 0
raw = subprocess.run(["ncu","-i",rep,"--page","raw","--csv"],capture_output=True,text=True).stdout
rows=list(csv.reader(io.StringIO(raw))); hdr=rows[0]; units=rows[1]; d=rows[2+launch]; raw_hdr=list(hdr)
keys=['Kernel Name','gpu__time_duration.sum','launch__grid_size','launch__registers_per_thread','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active','sm__throughput.avg.pct_of_peak_sustained_elapsed','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','dram__bytes_read.sum','dram__bytes_write.sum']
keys+=[h for h in hdr if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio')]
for k in keys:
    if k in hdr:
        i=hdr.index(k)
        try:
            if k.startswith('smsp__average') and float(d[i])<0.05: continue
        except: pass
        print(f"{k:90s} {d[i][:60]:>20s} {units[i]}")
src = subprocess.run(["ncu","-i",rep,"--page","source","--csv","--print-source","sass","--launch-skip",str(launch),"--launch-count","1"],capture_output=True,text=True).stdout
rows=list(csv.reader(io.StringIO(src)))
hdr=None; b=[]
for r in rows:
    if r and r[0]=="Address": hdr=r; continue
    if hdr and len(r)>6: b.append(r)
iE=hdr.index("Instructions Executed"); iS=hdr.index("# Samples")
ops=collections.Counter(); samp=collections.Counter(); tot=0
for r in b:
    t=r[1].split(); op=t[1] if t[0].startswith('@') else t[0]; op=op.split('.')[0]
    n=int(r[iE]); ops[op]+=n; samp[op]+=int(r[iS]); tot+=n
nw=nrays/32
# (a report with several launches repeats the table: normalise to the launch's own smsp__inst_executed.sum)
try:
    scale=float(d[raw_hdr.index('smsp__inst_executed.sum')].replace(',',''))/tot
except Exception:
    scale=1.0
nw/=scale
print("warp-instructions per ray: %.0f"%(tot/nw))
f64=sum(ops[o] for o in ('DFMA','DMUL','DADD','DSETP'))/nw
print("FP64-pipe instructions per ray: %.0f"%f64)
print(' '.join("%s %.0f"%(o,n/nw) for o,n in ops.most_common(24)))
