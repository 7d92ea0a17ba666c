#!/usr/bin/env python
"""Code footprint of a kernel from an .ncu-rep (--import-source on): SASS instructions in address order, how many of them
were executed at all, and the executed footprint in 128-byte instruction-cache lines; then a map of the address space in
chunks (default 4 KB): instructions executed, warp-level executions, no-instruction stall samples.

    python tools/ncu_code_footprint.py report.ncu-rep [chunk bytes]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
chunk = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = next(r for r in rows if r and r[0] == "Address")
ix = {n: i for i, n in enumerate(hdr)}
col_no_inst = next((i for i, n in enumerate(hdr) if "no_instruction" in n.lower() or "No Instruction" in n), None)
ins = []
for r in rows:
    if len(r) == len(hdr) and r[0].startswith("0x"):
        ins.append((int(r[0], 16), int(r[ix["Instructions Executed"]] or 0), int(r[ix["# Samples"]] or 0),
                    int(r[col_no_inst] or 0) if col_no_inst is not None else 0, r[1].strip()))
base = ins[0][0]
total = len(ins)
hot = [i for i in ins if i[1] > 0]
lines = {(a - base) // 128 for a, e, *_ in hot}
print(f"SASS instructions {total} ({total * 16 / 1024:.1f} KB); executed at least once {len(hot)} ({len(hot) * 16 / 1024:.1f} KB); "
      f"executed 128-byte lines {len(lines)} ({len(lines) * 128 / 1024:.1f} KB)")
ex_total = sum(i[1] for i in ins)
# footprint weighted: lines that hold 90 / 99 % of the executed instructions
per_line = {}
for a, e, *_ in ins:
    per_line[(a - base) // 128] = per_line.get((a - base) // 128, 0) + e
acc = 0
srt = sorted(per_line.values(), reverse=True)
for pct in (0.5, 0.9, 0.99):
    acc = 0
    for k, v in enumerate(srt):
        acc += v
        if acc >= pct * ex_total:
            print(f"{int(pct * 100)} % of executed warp-instructions come from {k + 1} lines ({(k + 1) * 128 / 1024:.1f} KB)")
            break
print(f"{'offset':>8} {'static':>6} {'hot':>5} {'executed':>12} {'share':>6} {'samples':>8} {'no_inst':>8}")
nchunks = (ins[-1][0] - base) // chunk + 1
for c in range(nchunks):
    sel = [i for i in ins if (i[0] - base) // chunk == c]
    ex = sum(i[1] for i in sel)
    print(f"{c * chunk:8d} {len(sel):6d} {sum(1 for i in sel if i[1] > 0):5d} {ex:12d} {ex / max(ex_total, 1):6.3f} {sum(i[2] for i in sel):8d} {sum(i[3] for i in sel):8d}")
