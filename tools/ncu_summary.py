#!/usr/bin/env python
"""Summarises an .ncu-rep (raw metrics + per-source-file instruction shares) into text for profiles/."""
import collections, csv, subprocess, sys, io, re

def main(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    want = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
            'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
            'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
            'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
            'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
            'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
            'smsp__sass_average_branch_targets_threads_uniform.pct', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
            'smsp__cycles_active.avg', 'sm__cycles_elapsed.max', 'smsp__warps_eligible.avg.per_cycle_active',
            'smsp__warps_active.avg.per_cycle_active', 'l1tex__t_sector_hit_rate.pct', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
            'launch__shared_mem_per_block_dynamic', 'launch__waves_per_multiprocessor', 'sm__maximum_warps_per_active_cycle_pct']
    print("kernel:", vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?")
    for i, h in enumerate(hdr):
        if h in want or re.match(r"smsp__average_warps_issue_stalled_.*_per_issue_active.ratio", h) or re.match(r"smsp__average_warp_latency_issue_stalled_.*\.ratio", h):
            try:
                v = float(vals[i].replace(",", ""))
            except ValueError:
                continue
            if 'stalled' in h and v < 0.3:
                continue
            print(f"  {h:90s} {units[i]:12s} {vals[i]}")
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    cur = None; hdr = None
    agg = collections.Counter(); aggT = collections.Counter(); perline = collections.Counter(); perlineT = collections.Counter()
    for r in rows:
        if len(r) >= 2 and r[0] == 'File Path': cur = r[1].split('/')[-1]; hdr = None; continue
        if r and r[0] == 'Line No': hdr = r; continue
        if hdr and cur and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            if d['Address'] in ('', '-'):
                try: ie = int(d['Instructions Executed']); te = int(d['Thread Instructions Executed'])
                except ValueError: continue
                agg[cur] += ie; aggT[cur] += te; perline[(cur, int(d['Line No']))] += ie; perlineT[(cur, int(d['Line No']))] += te
    tot = sum(agg.values())
    print("instructions executed by source file (inlined code attributed to its own file):")
    for k, v in agg.most_common():
        if v: print(f"  {k:28s} inst={v:12d} share={v/tot:.3f} avg_active_threads={aggT[k]/max(v,1):5.1f}")
    print("top source lines:")
    for (f, l), v in perline.most_common(25):
        print(f"  {f}:{l:<5d} inst={v:11d} share={v/tot:.3f} avg_active_threads={perlineT[(f,l)]/max(v,1):5.1f}")

if __name__ == "__main__":
    main(sys.argv[1])
