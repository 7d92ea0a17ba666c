#!/usr/bin/env python
"""GPU-box helper: find strikes whose results differ between prune mode 0 and 1 and print both event logs.
    python tools/find_prune_diff.py <n_balls> <strikes per table> <tables> <seed0>"""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from billiard_ai_b200 import Engine, workloads as W
from prune_audit import rand_layout

n, nc, tables, seed0 = (int(a) for a in sys.argv[1:5])
eng = Engine.from_table(W.default_table(), W.SNOOKER_SCORES)
found = []
for t in range(tables):
    rng = np.random.RandomState(seed0 + t)
    pos = rand_layout(rng, n)
    types = np.array([0] + [1] * (n - 1), dtype=np.int32)
    ang = rng.uniform(0, 2 * np.pi, nc); spd = rng.uniform(50, 5000, nc)
    vx = (spd * np.cos(ang)).astype(np.float32); vy = (spd * np.sin(ang)).astype(np.float32)
    eng.set_states(pos, types, 2)
    eng.set_prune(0)
    off = {k: v.copy() for k, v in eng.evaluate(0, vx, vy, 100000, 0, final_pos=False, event_log=True).items() if v is not None}
    eng.set_prune(1)
    on = eng.evaluate(0, vx, vy, 100000, 0, final_pos=False, event_log=True)
    d = np.nonzero((off["event_log"] != on["event_log"]).reshape(nc, -1).any(axis=1) | (off["cost"] != on["cost"]))[0]
    for i in d.tolist():
        rec = dict(seed=seed0 + t, strike=i, vx=float(vx[i]), vy=float(vy[i]), vx_bits=int(vx[i:i+1].view(np.uint32)[0]),
                   vy_bits=int(vy[i:i+1].view(np.uint32)[0]), pos=pos.tolist(),
                   log_off=off["event_log"][i].tolist(), log_on=on["event_log"][i].tolist())
        found.append(rec)
        print("DIFF seed", seed0 + t, "strike", i, "v", vx[i], vy[i], flush=True)
        for e in range(32):
            a, b = off["event_log"][i][e], on["event_log"][i][e]
            fa = np.array([a[3]], np.int32).view(np.float32)[0]; fb = np.array([b[3]], np.int32).view(np.float32)[0]
            print("  ", e, a[:3].tolist(), fa, "|", b[:3].tolist(), fb, "" if (a == b).all() else "  <--")
            if not (a == b).all():
                break
json.dump(found, open(os.path.join(ROOT, "gpurun_out", "prune_diff.json"), "w"))
print("found", len(found))
