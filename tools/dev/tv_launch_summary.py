#!/usr/bin/env python
"""Per-ADMM-iteration breakdown of the TV-1D kernels from an ncu launch list (gpu__time_duration.sum)."""
import csv, re, sys
path = sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/tv_launches.csv'
lines=[l for l in open(path) if not l.startswith('==')]
rows=[(x['Kernel Name'], float(x['Metric Value'].replace(',',''))/1000) for x in csv.DictReader(lines)]
def short(n):
    n=re.sub(r'\(.*','',n); n=re.sub(r'<.*','',n); return n.split('::')[-1][:24]
seq=[(short(n),v) for n,v in rows]
its=[];cur=[]
for n,v in seq:
    if (n.startswith('TvpPreSum') or n.startswith('TvPreKernel')) and cur: its.append(cur);cur=[]
    cur.append((n,v))
its.append(cur)
for it in its[1:]:
    tv=[(n,v) for n,v in it if n.startswith('Tvp')]
    print("iter: total %.0f us, tv %.0f us (%d launches) |"%(sum(v for n,v in it), sum(v for n,v in tv), len(tv)), ' '.join("%s:%.0f"%(n[3:8],v) for n,v in tv), '| other', ' '.join("%s:%.0f"%(n[:8],v) for n,v in it if not n.startswith('Tvp')))
