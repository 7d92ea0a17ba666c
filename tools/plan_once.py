#!/usr/bin/env python
"""GPU-box helper for ncu passes: the device planner on a beam of 1024 random 16-ball states, planned and simulated twice."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from billiard_ai_b200 import Engine, workloads as W   # noqa: E402

tbl = W.default_table()
eng = Engine.from_table(tbl, W.SNOOKER_SCORES, cue_type=W.TYPE_CODE["WHITE"])
eng.set_planner(tbl, velocity_multiply=True)
pos, types = W.random_tables(1024, 16, seed0=5000)
eng.set_states(pos, types, 1 << W.TYPE_CODE["RED"])
for _ in range(2):
    n = eng.plan()
    eng.evaluate_plan(final_pos=False)
print("planned", n)
