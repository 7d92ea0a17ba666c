#!/usr/bin/env python
"""ncu launch list (--csv) with the per-opcode thread-instruction counters -> profiles/rNN_level_fp_ops.json.

    ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,\
smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,\
smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,\
smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum \
        --clock-control none --csv --log-file gpurun_out/fp_ops.csv python tools/level_once.py 1
    python tools/ncu_fp_ops.py gpurun_out/fp_ops.csv gpurun_out/level_counters_prune1_n22.json "note" > profiles/r02_level_fp_ops.json

Per kernel of ONE level (the last init_kernel ... pack_records_kernel run of the list): launches, time, warp instructions,
thread instructions, DADD / DMUL / DFMA / FADD / FMUL / FFMA thread instructions; and the figure bench.py uses:
fp64_flop_per_quartic = (DADD + DMUL + 2 DFMA thread instructions of quartic_kernel) / quartics solved (the level's own
counter).  These are instructions EXECUTED (divisions, square roots and the libm restatement are made of them), which is
what the FP64 pipe sees; the `useful` arithmetic of the closed form is a small part of it."""
import collections
import csv
import json
import re
import sys

SHORT = {"smsp__sass_thread_inst_executed_op_dadd_pred_on.sum": "dadd", "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum": "dmul",
         "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum": "dfma", "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum": "fadd",
         "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum": "fmul", "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum": "ffma",
         "smsp__inst_executed.sum": "warp_inst", "smsp__thread_inst_executed.sum": "thread_inst", "gpu__time_duration.sum": "us",
         "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write"}


def main(path, counters_path, note):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr, rows = rows[0], rows[1:]
    ki, mi, ui, vi = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Unit"), hdr.index("Metric Value")
    launches = collections.OrderedDict()
    for r in rows:
        name = re.sub(r"\(.*", "", r[ki]).replace("unnamed>::", "").split("::")[-1]
        name = re.sub(r"_(dense|wide)$", "", name)   # the kernels for large parts: same work per item / job
        v = float(r[vi].replace(",", ""))
        if r[ui] == "ns":
            v /= 1e3
        elif r[ui] == "ms":
            v *= 1e3
        elif r[ui] in ("Kbyte", "Mbyte", "Gbyte"):
            v *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[r[ui]]
        launches.setdefault(int(r[0]), dict(kernel=name))[SHORT.get(r[mi], r[mi])] = v
    seq = list(launches.values())
    ends = [i for i, l in enumerate(seq) if l["kernel"] == "pack_records_kernel"]
    starts = [i for i, l in enumerate(seq) if l["kernel"] == "init_kernel"]
    flush = [i for i, l in enumerate(seq) if l["kernel"] == "flush_kernel"]
    end = max(e for e in ends if not flush or e < flush[-1])          # the warm level (no flush in front)
    # a level runs as several parts on concurrent streams; under ncu they are serialised part after part (init ... finish,
    # init ... finish, selection): the level starts where the run of pipeline kernels in front of the selection starts
    pipe = {"init_kernel", "items_kernel", "quartic_kernel", "apply_kernel", "finish_kernel", "counters_kernel"}
    start = end
    while start > 0 and not seq[start - 1]["kernel"] in pipe:
        start -= 1
    while start > 0 and seq[start - 1]["kernel"] in pipe:
        start -= 1
    level = seq[start:end + 1]
    per = collections.OrderedDict()
    for l in level:
        d = per.setdefault(l["kernel"], collections.defaultdict(float))
        d["launches"] += 1
        for k, v in l.items():
            if k != "kernel":
                d[k] += v
    cnt = json.load(open(counters_path))
    q = per["quartic_kernel"]
    fp64 = q["dadd"] + q["dmul"] + 2 * q["dfma"]
    tot = lambda m: sum(l.get(m, 0.0) for l in level)   # noqa: E731
    out = dict(source=note, workload=f"{cnt['n_balls']} balls, {cnt['strikes']} strikes, prune mode {cnt['prune']}", counters=cnt,
               launches=len(level), serialised_us=tot("us"), warp_instructions=tot("warp_inst"), thread_instructions=tot("thread_inst"),
               dram_bytes_read=tot("dram_read"), dram_bytes_write=tot("dram_write"),
               fp64_thread_inst=dict(dadd=tot("dadd"), dmul=tot("dmul"), dfma=tot("dfma")),
               fp32_thread_inst=dict(fadd=tot("fadd"), fmul=tot("fmul"), ffma=tot("ffma")),
               fp64_flop_per_quartic=fp64 / max(cnt["quartics"], 1),
               fp64_inst_per_quartic=(q["dadd"] + q["dmul"] + q["dfma"]) / max(cnt["quartics"], 1),
               thread_inst_per_quartic=q["thread_inst"] / max(cnt["quartics"], 1),
               per_kernel={k: dict(v) for k, v in per.items()})
    json.dump(out, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
