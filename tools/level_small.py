#!/usr/bin/env python
"""GPU-box helper: a SMALL level (default 1024 strikes, 22 balls) a few times -- the fixed cost of the 31-round chain."""
import os
import sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from billiard_ai_b200 import Engine, workloads as W   # noqa: E402
nd = int(sys.argv[1]) if len(sys.argv) > 1 else 32
eng = Engine.from_table(W.default_table(), W.SNOOKER_SCORES, cue_type=W.TYPE_CODE["WHITE"])
pos, types = W.snooker22_open()
dirs, speeds = W.grid(nd, 32)
eng.set_prune(1)
eng.set_states(pos, types, 1 << W.TYPE_CODE["RED"])
eng.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
ms = []
for _ in range(20):
    eng.rerun()
    ms.append(eng.last_kernel_ms())
print("strikes", nd * 32, "kernel ms median", float(np.median(ms)), "min", min(ms))
