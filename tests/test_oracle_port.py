"""The plain C++ restatement (oracle/port) must agree bit-for-bit with the reference build (oracle/_ref)."""
import os

import numpy as np

from oracle_lib import GOLDEN, TYPE_CODE


def _rand_layout(rng, n):
    pos = []
    while len(pos) < n:
        p = np.array([rng.uniform(-880, 880), rng.uniform(-420, 420)], dtype=np.float32)
        if all(np.hypot(*(p - q)) > 53.3 for q in pos):
            pos.append(p)
    return np.array(pos, dtype=np.float32)


def test_port_matches_committed_reference_vectors(port):
    g = np.load(os.path.join(GOLDEN, "ref_vectors.npz"))
    ev = port.evaluate(g["rand16_pos"][None], g["rand16_types"][None], 1 << 1, 0, g["rand16_vx"], g["rand16_vy"],
                       100000, 0, threads=4, event_log=True)
    clean = ev["min_gap"] >= 1e-9  # north_star tie clause
    assert clean.mean() > 0.95
    for k in ("flags", "n_events", "potted", "cost"):
        assert np.array_equal(ev[k][clean], g["rand16_" + k][clean]), k
    assert np.array_equal(ev["final_time"][clean].view(np.uint32), g["rand16_final_time"][clean].view(np.uint32))
    assert np.array_equal(ev["final_pos"][clean].view(np.uint32), g["rand16_final_pos"][clean].view(np.uint32))
    # event logs: type, balls (unordered pair for collisions), layer time bits
    a, b = ev["event_log"][clean], g["rand16_event_log"][clean]
    assert np.array_equal(a[..., 0], b[..., 0]) and np.array_equal(a[..., 3], b[..., 3])
    lo = np.minimum(b[..., 1], b[..., 2]); hi = np.maximum(b[..., 1], b[..., 2])
    coll = b[..., 0] == 0
    assert np.array_equal(np.where(coll, lo, b[..., 1]), a[..., 1])
    assert np.array_equal(np.where(coll, hi, b[..., 2]), a[..., 2])


def test_port_matches_reference_live(port, ref):
    rng = np.random.RandomState(7)
    for n, nc in ((2, 300), (6, 300), (22, 200)):
        pos = _rand_layout(rng, n)
        types = np.array([0] + [1] * (n - 1), dtype=np.int32)
        ang = rng.uniform(0, 2 * np.pi, nc); spd = rng.uniform(100, 5000, nc)
        vx = (spd * np.cos(ang)).astype(np.float32); vy = (spd * np.sin(ang)).astype(np.float32)
        sc = rng.randint(1, 10**6, nc).astype(np.uint64); bc = rng.randint(0, 10**6, nc).astype(np.uint64)
        a = port.evaluate(pos[None], types[None], 2, 0, vx, vy, sc, bc, threads=4)
        b = ref.evaluate(pos[None], types[None], 2, 0, vx, vy, sc, bc, threads=4)
        clean = a["min_gap"] >= 1e-9
        assert clean.mean() > 0.9
        for k in ("flags", "n_events", "potted", "cost"):
            assert np.array_equal(a[k][clean], b[k][clean]), (n, k)
        assert np.array_equal(a["final_pos"][clean].view(np.uint32), b["final_pos"][clean].view(np.uint32))
        assert np.array_equal(a["final_time"][clean].view(np.uint32), b["final_time"][clean].view(np.uint32))


def test_port_trace_matches_reference(port, ref):
    rng = np.random.RandomState(11)
    pos = _rand_layout(rng, 8)
    types = np.array([0, 1, 1, 1, 2, 3, 6, 7], dtype=np.int32)
    checked = 0
    for i in range(60):
        ang = rng.uniform(0, 2 * np.pi); spd = rng.uniform(500, 5000)
        vx, vy = np.float32(spd * np.cos(ang)), np.float32(spd * np.sin(ang))
        nl_p, t_p, rec_p, gap = port.simulate_trace(pos, types, vx, vy)
        nl_r, t_r, rec_r = ref.simulate_trace(pos, types, vx, vy)
        if gap < 1e-9:
            continue
        assert nl_p == nl_r
        assert np.array_equal(t_p.view(np.uint32), t_r.view(np.uint32))
        # slots 0..8 (kind, after pos/vel/rolling, before vel, cause); slot 9 is not tracked by the port
        assert np.array_equal(rec_p[..., :9].view(np.uint32), rec_r[..., :9].view(np.uint32))
        checked += 1
    assert checked > 40
