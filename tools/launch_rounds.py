#!/usr/bin/env python
"""Per-round kernel times of the first pipeline pass in an ncu launch list (gpu__time_duration.sum csv)."""
import csv, sys, collections
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>5]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); ui=hdr.index('Metric Unit')
seq=[]
for r in rows[1:]:
    name=r[ki].split('(')[0].split('::')[-1]
    v=float(r[vi].replace(',',''))
    if r[ui]=='ns': v/=1000
    elif r[ui]=='ms': v*=1000
    seq.append((name,v))
idx=[i for i,(n,v) in enumerate(seq) if 'init_kernel' in n]
a=idx[0]; b=idx[1] if len(idx)>1 else len(seq)
tot=collections.Counter()
for n,v in seq[a:b]: tot[n]+=v
print({k:round(v,1) for k,v in tot.items()}, 'sum us', round(sum(tot.values()),1))
i=a+1; r=0
while i+2<b and 'items' in seq[i][0]:
    print(r, [round(seq[i+k][1],1) for k in range(3)])
    i+=3; r+=1
