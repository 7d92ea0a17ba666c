#!/usr/bin/env python
"""Per-line instruction counts of one source file in an .ncu-rep: tools/ncu_ranges.py rep file.cuh [min_share]"""
import collections, csv, subprocess, sys, io
rep, fname = sys.argv[1], sys.argv[2]
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.003
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
cur = None; hdr = None
per = collections.Counter(); perT = collections.Counter(); perS = collections.Counter(); tot = 0; totS = 0
for r in rows:
    if len(r) >= 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; hdr = None; continue
    if r and r[0] == 'Line No': hdr = r; continue
    if hdr and cur and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        if d['Address'] in ('', '-'):
            try: ie = int(d['Instructions Executed']); te = int(d['Thread Instructions Executed'])
            except ValueError: continue
            try: sm = int(d.get('# Samples', '0') or 0)
            except ValueError: sm = 0
            tot += ie; totS += sm
            if cur == fname: per[int(d['Line No'])] += ie; perT[int(d['Line No'])] += te; perS[int(d['Line No'])] += sm
lines = open(subprocess.run(["find", ".", "-name", fname], capture_output=True, text=True).stdout.split()[0]).read().split("\n")
for l in sorted(per):
    if per[l] / tot >= thr or perS[l] / max(totS, 1) >= thr:
        print(f"{l:4d} inst {per[l]/tot*100:5.2f}% time {perS[l]/max(totS,1)*100:5.2f}% act {perT[l]/max(per[l],1):4.1f} | {lines[l-1].strip()[:100]}")
print("file share: inst %.1f%% time %.1f%%" % (sum(per.values()) / tot * 100, sum(perS.values()) / max(totS, 1) * 100))
