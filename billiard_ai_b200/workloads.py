"""Synthetic workloads of BASELINE.json `configs` as restated in SURVEY.md 8(d).

The reference has no such grids (its planner emits <100 candidates per search, billiard_search.cpp:797-826)
and no spin parameter; these are scale-ups of the per-candidate evaluation, which is what has reference
semantics.  Everything here is plain numpy on the host; identical buffers go to the engine and to the oracle.
"""
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# type codes used throughout the repo
TYPE_NAMES = ["WHITE", "RED", "YELLOW", "GREEN", "BROWN", "BLUE", "PINK", "BLACK"]
TYPE_CODE = {n: i for i, n in enumerate(TYPE_NAMES)}
# billiard_snooker.cpp:751-759 points, :787-790 score = points / 7.0
SNOOKER_SCORES = np.array([0.0, 1 / 7.0, 2 / 7.0, 3 / 7.0, 4 / 7.0, 5 / 7.0, 6 / 7.0, 7 / 7.0], dtype=np.float64)
GRID_SEARCH_COST = 100000  # 1.0 in the reference's cost units (COST_FLOAT_TO_INT_FACTOR, billiard_search.cpp:26)

# billiard_ui/Assets/Resources/configuration.json "spots"
SNOOKER_SPOTS = {"BLACK": (893.8, 0.0), "PINK": (580.0, 0.0), "BLUE": (0.0, 0.0), "BROWN": (-600.0, 0.0),
                 "GREEN": (-600.0, -200.0), "YELLOW": (-600.0, 200.0)}


def default_table():
    """Table of test/billiard_search/resources/configuration.json (committed copy: tests/golden/table.json)."""
    with open(os.path.join(ROOT, "tests", "golden", "table.json")) as f:
        return json.load(f)


def snooker22(radius=26.15, gap=1.0):
    """Config 2/4 layout: colours on their spots, 15 reds in a triangle behind PINK, cue ball in the D."""
    pos, types = [(-680.0, -60.0)], [TYPE_CODE["WHITE"]]
    pitch = 2 * radius + gap
    apex_x = SNOOKER_SPOTS["PINK"][0] + pitch
    for row in range(5):
        x = apex_x + row * pitch * np.sqrt(3.0) / 2.0
        for k in range(row + 1):
            pos.append((x, (k - row / 2.0) * pitch))
            types.append(TYPE_CODE["RED"])
    for name in ("YELLOW", "GREEN", "BROWN", "BLUE", "PINK", "BLACK"):
        pos.append(SNOOKER_SPOTS[name])
        types.append(TYPE_CODE[name])
    return np.array(pos, dtype=np.float32), np.array(types, dtype=np.int32)


def random_layout(n, seed, radius=26.15, half_len=940.5, half_wid=471.5, min_gap=1.0):
    """n balls, seeded, centre distance >= 2r + min_gap, >= r (+1 mm) from the cushion lines.  Ball 0 is WHITE."""
    rng = np.random.RandomState(seed)
    pos = []
    lim_x, lim_y = half_len - 2 * radius - 1.0, half_wid - 2 * radius - 1.0
    while len(pos) < n:
        p = np.array([rng.uniform(-lim_x, lim_x), rng.uniform(-lim_y, lim_y)], dtype=np.float32)
        if all(np.hypot(*(p.astype(np.float64) - q)) >= 2 * radius + min_gap for q in pos):
            pos.append(p)
    types = np.array([TYPE_CODE["WHITE"]] + [TYPE_CODE["RED"]] * (n - 1), dtype=np.int32)
    return np.array(pos, dtype=np.float32), types


def snooker22_open(seed=22):
    """Config 2/4 layout used by bench.py: the 22 snooker balls (WHITE, 15 RED, 6 colours) at seeded random
    non-overlapping positions (SURVEY.md 8d "seeded random non-overlapping (seed 22)") -- an open mid-frame table.
    The racked layout of snooker22() has no valid RED pot within one strike (every candidate is invalid), which
    makes it an edge case for the tests rather than a search workload."""
    pos, _ = random_layout(22, seed)
    types = np.array([TYPE_CODE["WHITE"]] + [TYPE_CODE["RED"]] * 15 +
                     [TYPE_CODE[n] for n in ("YELLOW", "GREEN", "BROWN", "BLUE", "PINK", "BLACK")], dtype=np.int32)
    return pos, types


def random_tables(n_tables, n_balls, seed0=1000):
    """Config 5: table i seeded seed0 + i."""
    pos = np.zeros((n_tables, n_balls, 2), np.float32)
    types = np.zeros((n_tables, n_balls), np.int32)
    for i in range(n_tables):
        pos[i], types[i] = random_layout(n_balls, seed0 + i)
    return pos, types


def grid(n_dirs, n_speeds, max_speed=5000.0):
    """dirs[k] = (cos, sin)(2 pi k / n_dirs) rounded to binary32; speeds[j] = (j+1) * max_speed / n_speeds."""
    th = 2.0 * np.pi * np.arange(n_dirs, dtype=np.float64) / n_dirs
    dirs = np.stack([np.cos(th), np.sin(th)], axis=1).astype(np.float32)
    speeds = ((np.arange(n_speeds, dtype=np.float64) + 1.0) * (max_speed / n_speeds)).astype(np.float32)
    return dirs, speeds


def grid_velocities(dirs, speeds):
    """Explicit (vx, vy) of one state's grid in candidate-id order (dir-major, speed-minor), computed as the
    kernel does: one binary32 multiply per component."""
    vx = (speeds[None, :] * dirs[:, 0:1]).astype(np.float32).reshape(-1)
    vy = (speeds[None, :] * dirs[:, 1:2]).astype(np.float32).reshape(-1)
    return vx, vy


def shard_range(total, rank, world):
    """Contiguous id range of `rank`."""
    per = (total + world - 1) // world
    lo = min(total, rank * per)
    return lo, min(total, lo + per)
