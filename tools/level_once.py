#!/usr/bin/env python
"""GPU-box helper for ncu passes: the bench workload (config 2: 22 balls, 1024 x 64 strikes, beam 1024), a few warm-up levels,
then one level without and one level with an L2 flush in front.  Prints the work counters of the level as JSON
(gpurun_out/level_counters_prune<mode>.json) so that ncu instruction counts can be divided by quartics solved.

    python tools/level_once.py [prune mode = 1] [n_balls = 22]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from billiard_ai_b200 import Engine, workloads as W   # noqa: E402

prune = int(sys.argv[1]) if len(sys.argv) > 1 else 1
n_balls = int(sys.argv[2]) if len(sys.argv) > 2 else 22
eng = Engine.from_table(W.default_table(), W.SNOOKER_SCORES, cue_type=W.TYPE_CODE["WHITE"])
pos, types = W.snooker22_open() if n_balls == 22 else W.random_layout(n_balls, n_balls)
dirs, speeds = W.grid(1024, 64)
eng.set_prune(prune)
eng.set_states(pos, types, 1 << W.TYPE_CODE["RED"])
eng.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
eng.enable_counters(True)
eng.rerun()
cnt = eng.counters()
eng.enable_counters(False)
for _ in range(3):
    eng.level_rerun(1024)
eng.sync()
eng.level_rerun(1024)       # warm level
eng.flush_l2()
eng.level_rerun(1024)       # the level bench.py times (right after the flush)
eng.sync()
cnt["strikes"] = 65536
cnt["prune"] = prune
cnt["n_balls"] = n_balls
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(cnt, open(os.path.join(ROOT, "gpurun_out", f"level_counters_prune{prune}_n{n_balls}.json"), "w"))
print(json.dumps(cnt))
