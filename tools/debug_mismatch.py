#!/usr/bin/env python
"""GPU-box helper: find candidates whose engine result differs from the port oracle and dump both layer traces."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib
from parity import compare_batch, summary
from billiard_ai_b200 import Engine, workloads as W

def rand_layout(rng, n):
    pos = []
    while len(pos) < n:
        p = np.array([rng.uniform(-880, 880), rng.uniform(-420, 420)], dtype=np.float32)
        if all(np.hypot(*(p - q)) > 53.3 for q in pos):
            pos.append(p)
    return np.array(pos, dtype=np.float32)

def main():
    n, nc, seed = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    tbl = W.default_table()
    eng = Engine.from_table(tbl, W.SNOOKER_SCORES)
    port = oracle_lib.PortOracle(tbl)
    rng = np.random.RandomState(seed)
    pos = rand_layout(rng, n)
    types = np.array([0] + [1] * (n - 1), dtype=np.int32)
    ang = rng.uniform(0, 2 * np.pi, nc); spd = rng.uniform(100, 5000, nc)
    vx = (spd * np.cos(ang)).astype(np.float32); vy = (spd * np.sin(ang)).astype(np.float32)
    eng.set_states(pos, types, 2)
    out = eng.evaluate(0, vx, vy, 100000, 0, final_pos=True, event_log=True)
    ora = port.evaluate(pos[None], types[None], 2, 0, vx, vy, 100000, 0, threads=os.cpu_count(), event_log=True)
    c = compare_batch(out, ora)
    print(summary(c))
    bad = np.nonzero(~c["ok"] | ~c["bit_exact"])[0]
    np.set_printoptions(precision=9, linewidth=200)
    print("pos =", repr(pos))
    for j in bad[:4]:
        print("=== candidate", j, "v =", float(vx[j]).hex(), float(vy[j]).hex(), "min_gap", ora["min_gap"][j])
        print("gpu flags", out["flags"][j], "nev", out["n_events"][j], "| port flags", ora["flags"][j], "nev", ora["n_events"][j])
        for l in range(32):
            a, b = out["event_log"][j, l], ora["event_log"][j, l]
            if a[0] < 0 and b[0] < 0: break
            ta, tb = np.array([a[3]], np.int32).view(np.float32)[0], np.array([b[3]], np.int32).view(np.float32)[0]
            print(f"  ev{l}: gpu {a[:3]} t={ta!r}   port {b[:3]} t={tb!r}", "" if (a == b).all() else "  <<<")
        nl, t, rec = eng.simulate_trace(pos, types, vx[j], vy[j])
        nlp, tp, recp, gap = port.simulate_trace(pos, types, vx[j], vy[j])
        for l in range(min(nl, nlp)):
            d = (rec[l].view(np.uint32) != recp[l].view(np.uint32))[:, :9]
            if d.any():
                print("  first differing layer", l, "time", t[l], tp[l])
                for b in np.nonzero(d.any(axis=1))[0]:
                    print("    ball", b, "gpu ", rec[l, b, :9]); print("    ball", b, "port", recp[l, b, :9])
                print("  previous layer state (port):")
                for b in range(n): print("    ", b, [float(x).hex() for x in recp[l - 1, b, :9]])
                break
        eng.set_states(pos, types, 2)

if __name__ == "__main__":
    main()
