#!/usr/bin/env python
"""GPU-box helper: ONE large launch (a slice of config 5: 1000 random 16-ball tables x 64 x 64 strikes = 4 096 000 strikes),
device-timed, both prune modes -- the issue-bound end of the workload range, where occupancy choices show that the
65 536-strike level (latency-bound) hides.  Honours BCU_LIB (variant builds) and the BCU_* launch knobs.

    python tools/big_launch.py [name] [tables = 1000]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from billiard_ai_b200 import Engine, workloads as W   # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "default"
nt = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
eng = Engine.from_table(W.default_table(), W.SNOOKER_SCORES, cue_type=W.TYPE_CODE["WHITE"])
tp, tt = W.random_tables(nt, 16, seed0=1000)
d5, s5 = W.grid(64, 64)
eng.set_states(tp, tt, 1 << W.TYPE_CODE["RED"])
eng.evaluate_grid(d5, s5, W.GRID_SEARCH_COST)
res = {"name": name, "strikes": nt * 4096}
for prune in (1, 0):
    eng.set_prune(prune)
    eng.rerun(); eng.rerun()
    ms = []
    for _ in range(3):
        eng.flush_l2()
        eng.timer_begin()
        eng.rerun()
        ms.append(eng.timer_end())
    res[f"ms_prune{prune}"] = float(np.mean(ms))
    res[f"strikes_per_sec_prune{prune}"] = nt * 4096 / (float(np.mean(ms)) * 1e-3)
print("%-14s prune1 %.1f M strikes/s  prune0 %.1f M strikes/s" % (name, res["strikes_per_sec_prune1"] / 1e6, res["strikes_per_sec_prune0"] / 1e6))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", f"big_launch_{name}.json"), "w"))
