#!/usr/bin/env python
"""GPU-box helper: strikes/s of the batch path vs batch size (config-2 layout, n_dirs x 64 speeds), kernel time from CUDA events."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from billiard_ai_b200 import Engine, workloads as W

def main():
    sizes = [int(x) for x in sys.argv[1:]] or [256, 1024, 4096, 16384]
    tbl = W.default_table()
    eng = Engine.from_table(tbl, W.SNOOKER_SCORES, cue_type=W.TYPE_CODE["WHITE"], device=0)
    pos, types = W.snooker22_open()
    eng.set_states(pos, types, 1 << W.TYPE_CODE["RED"])
    for nd in sizes:
        dirs, speeds = W.grid(nd, 64)
        eng.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
        for _ in range(2): eng.rerun()
        ms = []
        for _ in range(5):
            eng.rerun(); ms.append(eng.last_kernel_ms())
        m = float(np.median(ms))
        print(f"n={nd*64:8d}  kernel {m:8.3f} ms  {nd*64/m/1e3:8.2f} M strikes/s", flush=True)
    eng.close()

if __name__ == "__main__":
    main()
