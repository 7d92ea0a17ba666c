#!/usr/bin/env python
"""ANALYSIS ONLY: pruning statistics of the travel-distance bound on the CPU restatement (tools/prune_study.cpp)."""
import ctypes as C, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from billiard_ai_b200 import workloads as W
import oracle_lib as O
so = os.path.join(ROOT, "oracle", "_ref", "libprune_study.so")
subprocess.check_call(["g++", "-std=c++17", "-O2", "-DNDEBUG", "-ffp-contract=off", "-fPIC", "-pthread", "-w", "-shared",
                       os.path.join(ROOT, "tools", "prune_study.cpp"), "-o", so])
O.PORT_SO = so
po = O.PortOracle()
nballs = int(sys.argv[1]) if len(sys.argv) > 1 else 22
ncand = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
pos, types = (W.snooker22_open() if nballs == 22 else W.random_layout(nballs, 16))
dirs, speeds = W.grid(1024, 64)
vx, vy = W.grid_velocities(dirs, speeds)
rng = np.random.RandomState(1)
sel = rng.choice(len(vx), ncand, replace=False)
vx = np.ascontiguousarray(vx[sel]); vy = np.ascontiguousarray(vy[sel])
out = np.zeros(16)
po.lib.study_run(C.c_int(len(types)), O._ptr(np.ascontiguousarray(pos), C.c_float), O._ptr(np.ascontiguousarray(types), C.c_int32),
                 C.c_int64(ncand), O._ptr(vx, C.c_float), O._ptr(vy, C.c_float), O._ptr(out, C.c_double))
names = "steps jobs jobsStatic jobsDyn jobsHit prunedNP pass1 pass2 pass1b pass2b violNP viol2 viol2b maxRho winnerIsPair prunedPartial".split()
for k, v in zip(names, out): print(f"{k:14s} {v:.6g}")
print("jobs/step %.2f; single-pass keeps %.2f; 2-pass(strike) runs %.2f (+%.2f first); 2-pass(ball) runs %.2f (+%.2f first)" % (
    out[1]/out[0], (out[1]-out[5])/out[0], out[7]/out[0], out[6]/out[0], out[9]/out[0], out[8]/out[0]))
