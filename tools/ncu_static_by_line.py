#!/usr/bin/env python
"""Static SASS size per source function range from an .ncu-rep (source page, cuda+sass correlation): how many SASS
instructions each range owns, how many of them executed at least once, and their executed warp-instructions.
    python tools/ncu_static_by_line.py report.ncu-rep file:lo-hi:name ..."""
import csv, io, subprocess, sys, collections
rep = sys.argv[1]
ranges = []
for a in sys.argv[2:]:
    f, rg, name = a.split(':'); lo, hi = rg.split('-'); ranges.append((f, int(lo), int(hi), name))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
cur = None; hdr = None; line = None
st = collections.Counter(); hot = collections.Counter(); ex = collections.Counter()
for r in rows:
    if len(r) >= 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; hdr = None; continue
    if r and r[0] == 'Line No': hdr = r; continue
    if hdr and cur and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        if d.get('Address', '') in ('', '-'):
            try: line = int(d['Line No'])
            except ValueError: line = None
        elif line is not None:
            name = next((n for (rf, lo, hi, n) in ranges if rf == cur and lo <= line <= hi), cur + ':other')
            try: e = int(d['Instructions Executed'] or 0)
            except ValueError: e = 0
            st[name] += 1; ex[name] += e
            if e > 0: hot[name] += 1
print(f"{'range':28s} {'static':>7} {'KB':>6} {'hot':>6} {'hotKB':>6} {'executed':>12}")
for n, v in sorted(st.items(), key=lambda kv: -hot[kv[0]]):
    print(f"{n:28s} {v:7d} {v * 16 / 1024:6.1f} {hot[n]:6d} {hot[n] * 16 / 1024:6.1f} {ex[n]:12d}")
print("total", sum(st.values()), "hot", sum(hot.values()))
