#!/usr/bin/env python
"""GPU-box helper: find strikes on which a guard-variant build of the reach bound (billiard_ai_b200/build.py GUARD_VARIANTS:
`restall` = bound applied to every resting partner, `qrel0` = no q-cancellation guard, ...) changes the result, i.e. where
prune mode 1 of the variant differs from mode 0.  The strikes go to gpurun_out/guard_cases.json in the layout of
tests/golden/prune_regressions.json (minus the reference results, which tests/golden/make_prune_regressions.py --merge adds
in the build container).

    python tools/find_guard_cases.py <variant> <n_balls> <strikes per table> <tables> <seed0> [max_cases]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from billiard_ai_b200 import Engine, workloads as W   # noqa: E402
from billiard_ai_b200.build import variant_path       # noqa: E402
from prune_audit import rand_layout                   # noqa: E402


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32).tolist()


def main():
    variant = sys.argv[1]
    n, nc, tables, seed0 = (int(a) for a in sys.argv[2:6])
    max_cases = int(sys.argv[6]) if len(sys.argv) > 6 else 6
    eng = Engine.from_table(W.default_table(), W.SNOOKER_SCORES, lib_path=variant_path(variant))
    out_path = os.path.join(ROOT, "gpurun_out", "guard_cases.json")
    found = json.load(open(out_path)) if os.path.exists(out_path) else []
    have = sum(1 for c in found if variant in c["defeats"])
    types = np.array([0] + [1] * (n - 1), dtype=np.int32)
    for t in range(tables):
        if have >= max_cases:
            break
        rng = np.random.RandomState(seed0 + t)
        pos = rand_layout(rng, n)
        ang = rng.uniform(0, 2 * np.pi, nc); spd = rng.uniform(50, 5000, nc)
        vx = (spd * np.cos(ang)).astype(np.float32); vy = (spd * np.sin(ang)).astype(np.float32)
        eng.set_states(pos, types, 2)
        eng.set_prune(0)
        off = {k: v.copy() for k, v in eng.evaluate(0, vx, vy, 100000, 0, final_pos=False, event_log=True).items() if v is not None}
        eng.set_prune(1)
        on = eng.evaluate(0, vx, vy, 100000, 0, final_pos=False, event_log=True)
        d = np.nonzero((off["event_log"] != on["event_log"]).reshape(nc, -1).any(axis=1) | (off["cost"] != on["cost"]))[0]
        for i in d.tolist()[:2]:
            lo, hi = max(0, i - 2), min(nc, i + 3)
            found.append(dict(name=f"{variant}_seed{seed0 + t}_strike{i}", defeats=[variant], n_balls=n, pos_bits=bits(pos),
                              types=types.tolist(), vx_bits=bits(vx[lo:hi]), vy_bits=bits(vy[lo:hi]), focus=i - lo,
                              note=f"found by tools/find_guard_cases.py {variant} {n} {nc} (table seed {seed0 + t})"))
            have += 1
            print("CASE", found[-1]["name"], flush=True)
    os.makedirs(os.path.dirname(out_path), exist_ok=True)
    json.dump(found, open(out_path, "w"))
    print(f"{variant}: {have} cases in {out_path}")


if __name__ == "__main__":
    main()
