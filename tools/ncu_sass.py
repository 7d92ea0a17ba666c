#!/usr/bin/env python
"""Dumps the SASS of an .ncu-rep with per-instruction executed counts, average active threads and stall samples.
usage: ncu_sass.py report.ncu-rep [min_share]   -> prints instructions whose share of executed warp-instructions >= min_share"""
import csv, io, subprocess, sys

def load(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    ia, isrc, ismp, iex, ithr = (hdr.index(x) for x in ("Address", "Source", "# Samples", "Instructions Executed", "Avg. Threads Executed"))
    res = []
    for r in rows[2:]:
        if len(r) != len(hdr):
            continue
        res.append((r[ia], r[isrc].strip(), int(r[ismp] or 0), int(r[iex] or 0), float(r[ithr] or 0)))
    return res

if __name__ == "__main__":
    ins = load(sys.argv[1])
    tot = sum(i[3] for i in ins) or 1
    smp = sum(i[2] for i in ins) or 1
    print(f"# {len(ins)} SASS instructions, {tot} executed warp-instructions, {smp} samples")
    for n, (a, s, sm, ex, th) in enumerate(ins):
        print(f"{n:5d} ex={ex:9d} ({100*ex/tot:5.2f}%) smp={sm:6d} ({100*sm/smp:5.2f}%) thr={th:5.1f}  {s}")
