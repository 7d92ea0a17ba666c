#!/usr/bin/env python
"""GPU-box helper: audit of the reach bound (include/billiard_cuda.h, bcu_set_prune).

For `tables` random layouts of `n` balls and `nc` random strikes each: run the batch path with the bound off, on and
in audit mode; off and on must agree bit for bit on every output (cost, flags, potted, final time, final positions, event
log), and the audit must report 0 failures.  Prints one summary line per table and a total.

    python tools/prune_audit.py <n_balls> <strikes per table> <tables> [seed0]
"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from billiard_ai_b200 import Engine, workloads as W


def rand_layout(rng, n):
    pos = []
    while len(pos) < n:
        p = np.array([rng.uniform(-880, 880), rng.uniform(-420, 420)], dtype=np.float32)
        if all(np.hypot(*(p - q)) > 53.3 for q in pos):
            pos.append(p)
    return np.array(pos, dtype=np.float32)


def run(eng, mode, vx, vy, log):
    eng.set_prune(mode)
    return eng.evaluate(0, vx, vy, 100000, 0, final_pos=True, event_log=log)


def main():
    n, nc, tables = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    seed0 = int(sys.argv[4]) if len(sys.argv) > 4 else 1
    log = nc <= 262144
    eng = Engine.from_table(W.default_table(), W.SNOOKER_SCORES)
    eng.enable_counters(True)
    tot = dict(strikes=0, quartics_off=0, quartics_on=0, marked=0, failures=0, diff=0)
    t0 = time.time()
    for t in range(tables):
        rng = np.random.RandomState(seed0 + t)
        pos = rand_layout(rng, n)
        types = np.array([0] + [1] * (n - 1), dtype=np.int32)
        ang = rng.uniform(0, 2 * np.pi, nc); spd = rng.uniform(50, 5000, nc)
        vx = (spd * np.cos(ang)).astype(np.float32); vy = (spd * np.sin(ang)).astype(np.float32)
        eng.set_states(pos, types, 2)
        off = {k: v.copy() for k, v in run(eng, 0, vx, vy, log).items()}
        q_off = eng.counters()["quartics"]
        on = {k: v.copy() for k, v in run(eng, 1, vx, vy, log).items()}
        q_on = eng.counters()["quartics"]
        run(eng, 2, vx, vy, False)
        pc = eng.prune_counters()
        diff = 0
        for k in off:
            a, b = off[k], on[k]
            if a.dtype.kind == "f":
                a, b = a.view(np.uint32), b.view(np.uint32)
            d = (a != b)
            diff += int(d.reshape(len(vx), -1).any(axis=1).sum()) if d.ndim else int(d)
        tot["strikes"] += nc; tot["quartics_off"] += q_off; tot["quartics_on"] += q_on
        tot["marked"] += pc["skipped"]; tot["failures"] += pc["audit_failures"]; tot["diff"] += diff
        print(f"table {t}: quartics off {q_off} on {q_on} ({q_on / max(q_off, 1):.3f}) marked {pc['skipped']} "
              f"audit failures {pc['audit_failures']} differing strikes {diff}", flush=True)
    print("TOTAL", tot, "%.1f s" % (time.time() - t0))
    sys.exit(0 if tot["failures"] == 0 and tot["diff"] == 0 else 1)


if __name__ == "__main__":
    main()
