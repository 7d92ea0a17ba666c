#!/usr/bin/env python
"""Per source line (file:line) executed warp-instructions + avg active threads + stall samples from an .ncu-rep, grouped
by function ranges given as file:start-end:name on the command line (optional)."""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
cur = None; hdr = None
ex = collections.Counter(); th = collections.Counter(); smp = collections.Counter()
for r in rows:
    if len(r) >= 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; hdr = None; continue
    if r and r[0] == 'Line No': hdr = r; continue
    if hdr and cur and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        if d.get('Address', '') in ('', '-'):
            try:
                ie = int(d['Instructions Executed']); te = int(d['Thread Instructions Executed']); sm = int(d.get('# Samples', 0) or 0)
            except ValueError:
                continue
            k = (cur, int(d['Line No']))
            ex[k] += ie; th[k] += te; smp[k] += sm
tot = sum(ex.values()) or 1; stot = sum(smp.values()) or 1
ranges = []
for a in sys.argv[2:]:
    f, rg, name = a.split(':'); lo, hi = rg.split('-'); ranges.append((f, int(lo), int(hi), name))
if ranges:
    agg = collections.Counter(); aggt = collections.Counter(); aggs = collections.Counter()
    for (f, l), v in ex.items():
        name = next((n for (rf, lo, hi, n) in ranges if rf == f and lo <= l <= hi), f + ':other')
        agg[name] += v; aggt[name] += th[(f, l)]; aggs[name] += smp[(f, l)]
    for n, v in agg.most_common():
        print(f"{n:28s} inst={v:10d} share={v/tot:6.3f} samples={aggs[n]/stot:6.3f} thr={aggt[n]/max(v,1):5.1f}")
else:
    for k, v in ex.most_common(60):
        print(f"{k[0]}:{k[1]:<5d} inst={v:10d} share={v/tot:6.3f} samples={smp[k]/stot:6.3f} thr={th[k]/max(v,1):5.1f}")
