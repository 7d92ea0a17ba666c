#!/usr/bin/env python
"""Where the end-to-end step (bench.py's e2e) spends its time: wall clock per C-ABI call, bench workload."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from billiard_ai_b200 import Engine, workloads as W

BEAM = 1024
table = W.default_table()
pos, types = W.snooker22_open(22)
dirs, speeds = W.grid(1024, 64)
eng = Engine.from_table(table, W.SNOOKER_SCORES, cue_type=W.TYPE_CODE["WHITE"])
vx, vy = W.grid_velocities(dirs, speeds)
n = len(vx)
pin = lambda a: torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
keep = [pin(vx), pin(vy), pin(np.zeros(n, np.int32)), pin(np.full(n, W.GRID_SEARCH_COST, np.int64)), pin(np.zeros(n, np.int64))]
hvx, hvy, hst, hsc, hbc = [t.numpy() for t in keep]
hsc = hsc.view(np.uint64); hbc = hbc.view(np.uint64)
acc = {}


def timed(name, fn):
    t0 = time.perf_counter()
    r = fn()
    acc[name] = acc.get(name, 0.0) + time.perf_counter() - t0
    return r


def step():
    timed("set_states", lambda: eng.set_states(pos, types, 1 << W.TYPE_CODE["RED"]))
    timed("evaluate_level", lambda: eng.evaluate(hst, hvx, hvy, hsc, hbc, final_pos=False, event_log=False, reuse_buffers=True,
                                                 select_k=BEAM))
    timed("fetch_beam", lambda: eng.fetch_beam())


for _ in range(3):
    step()
acc.clear()
K = 20
t0 = time.perf_counter()
for _ in range(K):
    step()
tot = time.perf_counter() - t0
print(f"step {tot / K * 1e3:.3f} ms  kernel {eng.last_kernel_ms():.3f} ms")
for k, v in acc.items():
    print(f"  {k:14s} {v / K * 1e3:.3f} ms")
