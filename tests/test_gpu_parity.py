"""GPU parity tests: the CUDA engine, called through the C ABI (include/billiard_cuda.h), against the oracles.

Run on the B200 box:  python -m pytest tests -m gpu -x -q
Nothing here reads /root/reference: the oracles are the prebuilt oracle/_ref/*.so (the port can be rebuilt from
oracle/port on the box) and the committed fixtures under tests/golden/.
"""
import json
import os

import numpy as np
import pytest

from oracle_lib import GOLDEN, TYPE_CODE, SNOOKER_SCORES, F_COMPLETED, F_VALID
from parity import compare_batch, summary, rel_close

pytestmark = pytest.mark.gpu

# Fraction of tie-free candidates that must agree with the oracle under the north_star rules.
# 1.0 = every candidate.  See DESIGN.md "Parity status" for what the residual (if any) is.
MIN_PARITY = float(os.environ.get("BCU_MIN_PARITY", "1.0"))


@pytest.fixture(scope="module")
def engine(table):
    from billiard_ai_b200 import Engine
    e = Engine.from_table(table, SNOOKER_SCORES, cue_type=TYPE_CODE["WHITE"])
    yield e
    e.close()


def _rand_layout(rng, n):
    pos = []
    while len(pos) < n:
        p = np.array([rng.uniform(-880, 880), rng.uniform(-420, 420)], dtype=np.float32)
        if all(np.hypot(*(p - q)) > 53.3 for q in pos):
            pos.append(p)
    return np.array(pos, dtype=np.float32)


def test_native_library_is_loaded(engine):
    import billiard_ai_b200
    assert os.path.exists(billiard_ai_b200.LIB_PATH)
    with open("/proc/self/maps") as f:
        assert "libbilliard_cuda.so" in f.read()
    assert b"sm_100a" in engine.lib.bcu_version()
    assert engine.get_prune() == 0   # the C ABI's default: every quartic the reference solves is solved


def test_host_libm_selects_the_fma_variants_this_build_restates():
    """The device libm (csrc/bcu_libm.cuh) restates glibc 2.39's x86-64 `_fma` ifunc variants; the oracles run on the
    host's glibc.  Both claims are checked on THIS box: the CPU must offer FMA + AVX2 (what the ifunc resolvers test), and
    the host-compiled copy of the device headers must reproduce the host's libm bit for bit (exp, log, pow, sincos, atan2,
    log1p; 200 000 arguments each) -- tests/test_libm_exact.py, which otherwise only runs in the build container."""
    with open("/proc/cpuinfo") as f:
        flags = next(ln for ln in f if ln.startswith("flags")).split()
    assert "fma" in flags and "avx2" in flags, "glibc would select non-FMA libm variants on this host: parity unpinned here"
    import test_libm_exact as T
    T.check_libm_functions(200_000)


def test_rail_geometry_matches_reference_ctor(engine, table):
    # golden railsDerived were produced by the reference's Rail ctor (type.cpp:56-72) via oracle/_ref
    want = np.array(table["railsDerived"], dtype=np.float32)
    got = engine.rail_derived()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def _shots():
    with open(os.path.join(GOLDEN, "filmed_shots.json")) as f:
        return json.load(f)["tests"]


@pytest.mark.parametrize("shot", [s for s in _shots() if s["enabled"]], ids=lambda s: s["name"])
def test_filmed_shot_structure_and_trace(engine, port, shot):
    """Layer counts / node kinds asserted by test/billiard_search/src/simulation_vs_reality.cpp, and the full
    per-layer trace against the oracle."""
    pos = np.array([b["position"] for b in shot["balls"]], dtype=np.float32)
    types = np.array([TYPE_CODE[b["type"]] for b in shot["balls"]], dtype=np.int32)
    nl, times, rec = engine.simulate_trace(pos, types, *shot["velocity"])
    if shot["expectedLayers"]:
        assert nl in shot["expectedLayers"]
    var_index = {b["var"]: i for i, b in enumerate(shot["balls"])}
    for e in shot["expects"]:
        if e.get("onlyIfLayers", nl) != nl or e.get("notIfLayers", -1) == nl or e["layer"] >= nl:
            continue
        r = rec[e["layer"], var_index[e["ball"]]]
        assert int(r[0]) == e["kind"], (shot["name"], e)
        if "afterRolling" in e:
            assert bool(r[5]) == e["afterRolling"]
    nl_p, t_p, rec_p, gap = port.simulate_trace(pos, types, *shot["velocity"])
    if gap >= 1e-9:
        assert nl == nl_p
        assert rel_close(times, t_p).all()
        assert np.array_equal(rec[..., 0], rec_p[..., 0]) and np.array_equal(rec[..., 8], rec_p[..., 8])
        assert rel_close(rec[..., 1:3], rec_p[..., 1:3], scale=1881.0).all()
        vs = np.maximum(np.abs(rec_p[..., 3:5]).max(), 1.0)
        assert rel_close(rec[..., 3:5], rec_p[..., 3:5], scale=vs).all()
        assert rel_close(rec[..., 6:8], rec_p[..., 6:8], scale=vs).all()


def test_committed_reference_vectors(engine, port):
    """tests/golden/ref_vectors.npz rand16_*: outputs of the reference itself (oracle/_ref) on seeded inputs.  The golden
    file has no tie information; the port oracle (bit-identical to the reference build) supplies it: every strike clear of
    near ties must agree, no slack."""
    g = np.load(os.path.join(GOLDEN, "ref_vectors.npz"))
    engine.set_states(g["rand16_pos"], g["rand16_types"], 1 << TYPE_CODE["RED"])
    out = engine.evaluate(0, g["rand16_vx"], g["rand16_vy"], 100000, 0, final_pos=True, event_log=True)
    ora = {k: g["rand16_" + k] for k in ("cost", "flags", "n_events", "potted", "final_time", "final_pos", "event_log")}
    tie = port.evaluate(g["rand16_pos"][None], g["rand16_types"][None], 1 << TYPE_CODE["RED"], 0, g["rand16_vx"], g["rand16_vy"],
                        100000, 0, threads=4)
    ora["min_gap"] = tie["min_gap"]
    c = compare_batch(out, ora)
    print(summary(c))
    assert c["n_clean"] >= 0.95 * c["n"]
    assert c["n_bit"] == c["n_clean"], summary(c)


@pytest.mark.parametrize("name", ["test_01", "situation1", "situation2"])
@pytest.mark.parametrize("vm", [0, 1])
def test_reference_planner_candidates_and_thesis_costs(engine, name, vm):
    """Config 1: the reference planner's own candidate list (golden, from oracle/_ref ref_plan) evaluated on the
    GPU must give the reference's solution costs -- incl. the totals printed in the thesis (72206, ...)."""
    g = np.load(os.path.join(GOLDEN, "ref_vectors.npz"))
    with open(os.path.join(GOLDEN, "layouts.json")) as f:
        lay = json.load(f)[name]
    pos = np.array([b["position"] for b in lay["balls"]], dtype=np.float32)
    types = np.array([TYPE_CODE[b["type"]] for b in lay["balls"]], dtype=np.int32)
    p = f"plan_{name}_vm{vm}_"
    expected = 0
    for t in lay.get("searchTypes", ["RED"]):
        expected |= 1 << TYPE_CODE[t]
    engine.set_states(pos, types, expected)
    out = engine.evaluate(0, g[p + "vx"], g[p + "vy"], g[p + "search_cost"], g[p + "base_cost"], final_pos=False)
    none = np.uint64(2**64 - 1)
    want = g[p + "result_cost"]
    is_solution = (out["flags"] & (F_COMPLETED | F_VALID)) == (F_COMPLETED | F_VALID)
    got = np.where(is_solution, out["cost"], none)
    assert np.array_equal(got, want)
    if name == "situation1" and vm == 0:
        with open(os.path.join(GOLDEN, "thesis_costs.json")) as f:
            thesis = json.load(f)
        ranked = sorted(int(c) for c in got if c != none)
        assert ranked[:10] == thesis["lists"]["s1_with_banks"]
        assert ranked[0] == 72206


@pytest.mark.parametrize("n,nc,seed", [(2, 2048, 3), (6, 4096, 5), (16, 8192, 16), (22, 4096, 22)])
def test_random_strikes_match_port_oracle(engine, port, n, nc, seed):
    rng = np.random.RandomState(seed)
    pos = _rand_layout(rng, n)
    types = np.array([0] + [1] * (n - 1), dtype=np.int32)
    ang = rng.uniform(0, 2 * np.pi, nc)
    spd = rng.uniform(100, 5000, nc)
    vx = (spd * np.cos(ang)).astype(np.float32)
    vy = (spd * np.sin(ang)).astype(np.float32)
    sc = rng.randint(1, 10**6, nc).astype(np.uint64)
    bc = rng.randint(0, 10**6, nc).astype(np.uint64)
    engine.set_states(pos, types, 1 << 1)
    out = engine.evaluate(0, vx, vy, sc, bc, final_pos=True, event_log=True)
    ora = port.evaluate(pos[None], types[None], 1 << 1, 0, vx, vy, sc, bc, threads=os.cpu_count() or 4, event_log=True)
    c = compare_batch(out, ora)
    print(f"N={n}: " + summary(c))
    assert c["n_clean"] > 0.9 * nc
    assert c["n_ok"] >= MIN_PARITY * c["n_clean"], summary(c)


def test_random_strikes_match_reference_build(engine, ref, port):
    """Same as above against the reference's own code (oracle/_ref/libref_oracle.so, prebuilt)."""
    rng = np.random.RandomState(101)
    n, nc = 16, 1024
    pos = _rand_layout(rng, n)
    types = np.array([0] + [1] * 9 + [2, 3, 4, 5, 6, 7], dtype=np.int32)
    ang = rng.uniform(0, 2 * np.pi, nc)
    spd = rng.uniform(100, 5000, nc)
    vx = (spd * np.cos(ang)).astype(np.float32)
    vy = (spd * np.sin(ang)).astype(np.float32)
    mask = (1 << 1) | (1 << 7)
    engine.set_states(pos, types, mask)
    out = engine.evaluate(0, vx, vy, 100000, 5, final_pos=True, event_log=False)
    ora = ref.evaluate(pos[None], types[None], mask, 0, vx, vy, 100000, 5, threads=os.cpu_count() or 4)
    # the reference build has no tie output; the port (bit-identical to it) has: mask with it and demand every strike
    ora["min_gap"] = port.evaluate(pos[None], types[None], mask, 0, vx, vy, 100000, 5, threads=os.cpu_count() or 4)["min_gap"]
    c = compare_batch(out, ora)
    print(summary(c))
    assert c["n_clean"] >= 0.95 * c["n"]
    assert c["n_bit"] == c["n_clean"], summary(c)


def test_multi_state_absent_balls_and_empty_batch(engine, port):
    """Ragged input: states with absent balls (type -1, as after pots), several states per launch, n = 0."""
    rng = np.random.RandomState(9)
    n, ns, per = 12, 5, 400
    pos = np.stack([_rand_layout(rng, n) for _ in range(ns)])
    types = np.tile(np.array([0] + [1] * (n - 1), dtype=np.int32), (ns, 1))
    for s in range(ns):
        types[s, rng.choice(np.arange(1, n), size=s, replace=False)] = -1
    masks = np.array([2, 2, 6, 2, 2], dtype=np.uint32)
    nc = ns * per
    cs = np.repeat(np.arange(ns, dtype=np.int32), per)
    ang = rng.uniform(0, 2 * np.pi, nc)
    spd = rng.uniform(100, 5000, nc)
    vx = (spd * np.cos(ang)).astype(np.float32)
    vy = (spd * np.sin(ang)).astype(np.float32)
    engine.set_states(pos, types, masks)
    out = engine.evaluate(cs, vx, vy, 100000, 0, final_pos=True, event_log=True)
    ora = port.evaluate(pos, types, masks, cs, vx, vy, 100000, 0, threads=4, event_log=True)
    c = compare_batch(out, ora)
    print(summary(c))
    assert c["n_ok"] >= MIN_PARITY * c["n_clean"], summary(c)
    empty = engine.evaluate(0, np.zeros(0, np.float32), np.zeros(0, np.float32), 1, 0)
    assert len(empty["cost"]) == 0


def test_grid_topk_gather_against_numpy_and_oracle(engine, port):
    """Grid decode (K3), block top-k (K2) and gather: ids, ordering and ties exactly as numpy computes them
    from the fetched per-candidate results; a sample of the grid against the oracle."""
    from billiard_ai_b200 import workloads as W
    pos, types = W.random_layout(16, 16)
    dirs, speeds = W.grid(256, 16)
    ns = 3
    spos = np.stack([pos, pos[::-1].copy(), pos])
    stypes = np.stack([types, types[::-1].copy(), types])
    base = np.array([0, 1000, 7], dtype=np.uint64)
    engine.set_states(spos, stypes, 1 << 1, base)
    total = engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
    assert total == ns * 256 * 16
    res = engine.fetch()
    vx, vy = W.grid_velocities(dirs, speeds)
    cs = np.repeat(np.arange(ns, dtype=np.int32), len(vx))
    ora = port.evaluate(spos, stypes, 1 << 1, cs, np.tile(vx, ns), np.tile(vy, ns), W.GRID_SEARCH_COST, base[cs], threads=4)
    c = compare_batch(res, ora)
    print(summary(c))
    assert c["n_ok"] >= MIN_PARITY * c["n_clean"], summary(c)
    # top-k == stable sort of (cost, id) over valid candidates
    valid = (res["flags"] & F_VALID) != 0
    ids = np.nonzero(valid)[0]
    order = np.lexsort((ids, res["cost"][ids]))
    for k in (1, 7, 256, 1024):
        tc, ti = engine.topk(k)
        kk = min(k, len(ids))
        assert np.array_equal(ti, ids[order][:kk].astype(np.uint64))
        assert np.array_equal(tc, res["cost"][ids][order][:kk])
    tc, ti = engine.topk(64)
    gpos, gpot, gt = engine.gather(ti)
    assert np.array_equal(gpos, res["final_pos"][ti.astype(np.int64)])
    assert np.array_equal(gpot, res["potted"][ti.astype(np.int64)])
    assert np.array_equal(gt, res["final_time"][ti.astype(np.int64)])
    # a shard of the grid gives the same per-candidate results and global ids
    lo, hi = W.shard_range(total, 1, 4)
    engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST, first=lo, count=hi - lo)
    part = engine.fetch()
    for key in ("cost", "flags", "potted", "n_events"):
        assert np.array_equal(part[key], res[key][lo:hi])
    tc2, ti2 = engine.topk(16)
    sel = ids[(ids >= lo) & (ids < hi)]
    o2 = np.lexsort((sel, res["cost"][sel]))
    assert np.array_equal(ti2, sel[o2][:16].astype(np.uint64))


def test_full_size_config2_properties(engine):
    """BASELINE config 2 at full size (22-ball snooker layout, 1024 x 64 = 65 536 strikes): size-independent
    properties -- determinism across launches, every candidate accounted for, flag consistency, sorted top-k."""
    from billiard_ai_b200 import workloads as W
    pos, types = W.snooker22_open()
    dirs, speeds = W.grid(1024, 64)
    engine.set_states(pos, types, 1 << 1)
    n = engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
    assert n == 65536
    a = engine.fetch()
    engine.rerun()
    b = engine.fetch()
    for key in ("cost", "flags", "potted", "n_events", "final_time", "final_pos"):
        assert np.array_equal(a[key], b[key]), key
    fl = a["flags"]
    done = (fl & F_COMPLETED) != 0
    assert np.all(((fl & 4) != 0) | ((fl & 8) != 0) | done)           # every candidate ends in exactly one way
    assert np.all(((fl & F_VALID) == 0) | done)                         # valid implies completed
    assert np.all((a["cost"] == np.uint64(2**64 - 1)) == ~done)
    assert np.all(a["n_events"][(fl & 4) != 0] == 30)                   # overflow = MAX_EVENTS reached
    assert np.all(a["n_events"][done] >= 1)
    assert np.all(a["cost"][done] <= W.GRID_SEARCH_COST)                # min(total, 1) * searchCost
    tc, ti = engine.topk(1024)
    assert len(tc) == 1024
    assert np.all(np.diff(tc.astype(np.int64)) >= 0)
    assert len(np.unique(ti)) == len(ti)
    # the device-resident level (top-k -> record block -> merge) returns the same beam
    engine.level_select(1024)
    beam = engine.fetch_beam()
    assert np.array_equal(beam["cost"], tc) and np.array_equal(beam["id"], ti)
    assert np.array_equal(beam["final_pos"], a["final_pos"][ti.astype(np.int64)])
    assert np.array_equal(beam["potted"], a["potted"][ti.astype(np.int64)])


def test_racked_layout_has_no_valid_strike_and_gives_an_empty_beam(engine, port):
    """Edge case: on the racked 22-ball layout no single strike pots a RED cleanly, so every candidate is invalid,
    the top-k is empty and the beam has zero entries (the reference's search would return no solution)."""
    from billiard_ai_b200 import workloads as W
    pos, types = W.snooker22()
    dirs, speeds = W.grid(128, 16)
    engine.set_states(pos, types, 1 << 1)
    engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
    res = engine.fetch()
    vx, vy = W.grid_velocities(dirs, speeds)
    ora = port.evaluate(pos[None], types[None], 1 << 1, 0, vx, vy, W.GRID_SEARCH_COST, 0, threads=4)
    c = compare_batch(res, ora)
    print(summary(c))
    assert c["n_ok"] >= MIN_PARITY * c["n_clean"], summary(c)
    assert not np.any(res["flags"] & F_VALID)
    tc, ti = engine.topk(64)
    assert len(tc) == 0
    engine.level_select(64)
    assert len(engine.fetch_beam()["cost"]) == 0


def test_cpp_host_mirror_returns_the_reference_system(ref):
    """billiard_ai_b200/host/billiard_cuda_search.hpp, compiled against the reference's own headers
    (oracle/_ref/libmirror_check.so): Engine::simulate(state, velocity) must return the same node::System as
    billiard::search::simulate(state, velocity, config) -- every node, BallState bit, cause id and the
    _nextRolling / _lastRailCollisions bookkeeping -- and nullopt for the same strikes."""
    import ctypes as C
    from oracle_lib import ORACLE_DIR, TYPE_NAMES
    so = os.path.join(ORACLE_DIR, "_ref", "libmirror_check.so")
    if not os.path.exists(so):
        pytest.skip("libmirror_check.so not built (needs /root/reference at build time)")
    lib = C.CDLL(so)
    rng = np.random.RandomState(77)
    total_compared = 0
    for n in (3, 9, 16):
        pos = _rand_layout(rng, n)
        types = np.array([0] + [1] * (n - 1), dtype=np.int32)
        types[-1] = 7
        ns = 150
        ang = rng.uniform(0, 2 * np.pi, ns)
        spd = rng.uniform(300, 5000, ns)
        vx = (spd * np.cos(ang)).astype(np.float32)
        vy = (spd * np.sin(ang)).astype(np.float32)
        names = (C.c_char_p * len(TYPE_NAMES))(*[t.encode() for t in TYPE_NAMES])
        compared, nullopt = C.c_int(0), C.c_int(0)
        bad = lib.mirror_check(C.c_int(n), pos.ctypes.data_as(C.POINTER(C.c_float)), names, C.c_int(len(TYPE_NAMES)),
                               types.ctypes.data_as(C.POINTER(C.c_int32)), C.c_int(ns),
                               vx.ctypes.data_as(C.POINTER(C.c_float)), vy.ctypes.data_as(C.POINTER(C.c_float)),
                               C.byref(compared), C.byref(nullopt))
        print(f"N={n}: {compared.value} systems compared, {nullopt.value} nullopt on both sides, {bad} differ")
        assert bad == 0
        total_compared += compared.value
    assert total_compared > 350


def _fresh_engine(table, **env):
    """An engine created under the given BCU_* environment (the knobs are read by bcu_create)."""
    from billiard_ai_b200 import Engine
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        return Engine.from_table(table, SNOOKER_SCORES, cue_type=TYPE_CODE["WHITE"])
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


RESULT_KEYS = ("cost", "flags", "potted", "n_events", "final_time", "final_pos")


def test_pipeline_knobs_do_not_change_results(engine, table):
    """The batch pipeline cuts a launch into parts on concurrent streams, captures it in a CUDA graph and sizes its
    quartic job queue for the common case (the host doubles it and re-runs the launch when a round overflows it).
    None of that may change a bit of the results: one stream / no graph / a one-job-per-strike queue (which overflows
    and is grown several times) against the default engine, event logs included."""
    from billiard_ai_b200 import workloads as W
    pos, types = W.snooker22_open()
    dirs, speeds = W.grid(256, 32)
    vx, vy = W.grid_velocities(dirs, speeds)
    engine.set_states(pos, types, 1 << 1)
    want = engine.evaluate(0, vx, vy, W.GRID_SEARCH_COST, 0, final_pos=True, event_log=True)
    for env in ({"BCU_STREAMS": 1}, {"BCU_STREAMS": 3, "BCU_NO_GRAPH": 1}, {"BCU_JOBS_PER_STRIKE": 1},
                {"BCU_PDL": 0}, {"BCU_PDL": 2, "BCU_STREAMS": 4}, {"BCU_PRUNE": 1}):
        e = _fresh_engine(table, **env)
        try:
            e.set_states(pos, types, 1 << 1)
            got = e.evaluate(0, vx, vy, W.GRID_SEARCH_COST, 0, final_pos=True, event_log=True)
            for key in RESULT_KEYS + ("event_log",):
                assert np.array_equal(got[key], want[key]), (env, key)
            again = e.evaluate(0, vx, vy, W.GRID_SEARCH_COST, 0, final_pos=True, event_log=True, reuse_buffers=True)
            for key in RESULT_KEYS + ("event_log",):   # page-locked, reused result buffers hold the same values
                assert np.array_equal(again[key], want[key]), (env, "pinned", key)
        finally:
            e.close()


def test_reach_bound_and_filters_do_not_change_results(engine):
    """The reach bound on ball pairs and the division-free rail / partner filters (bcu_items.cuh, bcu_set_prune) only
    skip work whose result cannot be the next event.  Mode 0 makes every test and solves every quartic the reference
    does; mode 1 (default) must give the same bits on every output, event logs included, and the audit mode (everything
    solved, skipped jobs marked and checked against the strike's best time) must report no failure.  Random 16- and
    22-ball tables (the moving-vs-resting noise regime of the closed form shows up about once per 5e5 such pairs, so
    the sample is sized to contain it: ~1e8 quartics)."""
    from billiard_ai_b200 import workloads as W
    rng = np.random.RandomState(2024)
    total_marked = 0
    try:
        engine.enable_counters(True)
        for t in range(12):
            n = 16 if t % 2 == 0 else 22
            pos, types = W.random_layout(n, 7000 + t)
            nc = 65536
            ang = rng.uniform(0, 2 * np.pi, nc); spd = rng.uniform(50, 5000, nc)
            vx = (spd * np.cos(ang)).astype(np.float32); vy = (spd * np.sin(ang)).astype(np.float32)
            engine.set_states(pos, types, 1 << 1)
            engine.set_prune(0)
            off = {k: v.copy() for k, v in engine.evaluate(0, vx, vy, W.GRID_SEARCH_COST, 0, final_pos=True, event_log=True).items()}
            q_off = engine.counters()["quartics"]
            engine.set_prune(1)
            on = engine.evaluate(0, vx, vy, W.GRID_SEARCH_COST, 0, final_pos=True, event_log=True)
            q_on = engine.counters()["quartics"]
            for key in RESULT_KEYS + ("event_log",):
                assert np.array_equal(off[key].view(np.uint32) if off[key].dtype == np.float32 else off[key],
                                      on[key].view(np.uint32) if on[key].dtype == np.float32 else on[key]), (t, key)
            assert q_on < q_off
            engine.set_prune(2)
            engine.evaluate(0, vx, vy, W.GRID_SEARCH_COST, 0, final_pos=False)
            pc = engine.prune_counters()
            assert pc["audit_failures"] == 0, (t, pc)
            assert pc["skipped"] > 0   # (the exact count depends on which offers a strike's key already held when read)
            total_marked += pc["skipped"]
        assert total_marked > 10_000_000
    finally:
        engine.set_prune(0)
        engine.enable_counters(False)


def test_launch_larger_than_one_pipeline_pass(engine):
    """More candidates than one pass takes (4 streams x 2^20 strikes): the launch is processed in several waves over the
    same workspace.  Every candidate must come out exactly as when its shard of the grid is evaluated on its own."""
    from billiard_ai_b200 import workloads as W
    pos, types = W.random_layout(6, 5)
    dirs, speeds = W.grid(8192, 64)
    ns = 9                                       # 9 x 8192 x 64 = 4 718 592 candidates > 4 x 2^20
    spos = np.stack([np.roll(pos, i, axis=0) for i in range(ns)])
    stypes = np.stack([types] * ns)
    stypes[:, :] = 1
    for i in range(ns):
        stypes[i, i % 6] = 0                     # the cue ball sits somewhere else in every state
    engine.set_states(spos, stypes, 1 << 1)
    total = engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
    assert total == ns * 8192 * 64
    rng = np.random.RandomState(3)
    windows = [(0, 4096), (2**20 - 2048, 4096), (4 * 2**20 - 1000, 3000), (total - 5000, 5000)]
    windows += [(int(o), 2048) for o in rng.randint(0, total - 2048, 4)]
    full = {off: engine.fetch(off, n) for off, n in windows}
    for off, n in windows:
        engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST, first=off, count=n)
        part = engine.fetch()
        for key in RESULT_KEYS:
            assert np.array_equal(part[key], full[off][key]), (off, key)


def test_topk_dense_level_takes_the_streaming_sorter(engine):
    """The selection has two device paths (bcu_topk.cu): rank-by-counting on the compacted list of valid strikes, and the
    streaming sorter + rank merge when more than 16384 strikes are valid.  A 786k-strike level (48 states) has enough
    valid strikes for the second; a 16k-strike slice of it runs the first.  Both must equal numpy's stable (cost, id) order."""
    from billiard_ai_b200 import workloads as W
    layouts = [W.random_layout(16, 16 + s) for s in range(48)]
    spos = np.stack([p for p, _ in layouts]); stypes = np.stack([t for _, t in layouts])
    base = (np.arange(48, dtype=np.uint64) * 3) % 11
    engine.set_states(spos, stypes, 0xFE, base)            # any object ball counts: many valid strikes
    dirs, speeds = W.grid(256, 64)
    total = engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
    res = engine.fetch(final_pos=False)
    ids = np.nonzero((res["flags"] & F_VALID) != 0)[0]
    assert len(ids) > 16384, len(ids)
    order = ids[np.lexsort((ids, res["cost"][ids]))]
    for k in (1, 100, 1024):
        tc, ti = engine.topk(k)
        assert np.array_equal(ti, order[:k].astype(np.uint64)) and np.array_equal(tc, res["cost"][order[:k]])
    lo, n = 300000, 16384
    engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST, first=lo, count=n)
    sel = ids[(ids >= lo) & (ids < lo + n)]
    assert 0 < len(sel) <= 16384
    o2 = sel[np.lexsort((sel, res["cost"][sel]))]
    tc, ti = engine.topk(1024)
    kk = min(1024, len(sel))
    assert np.array_equal(ti, o2[:kk].astype(np.uint64)) and np.array_equal(tc, res["cost"][o2[:kk]])


def test_evaluate_level_equals_evaluate_then_select(engine):
    """bcu_evaluate_level (selection queued behind the simulation, result copies on a second stream) returns the same
    per-strike results and the same beam as bcu_evaluate followed by bcu_level_select."""
    from billiard_ai_b200 import workloads as W
    pos, types = W.snooker22_open(22)
    rng = np.random.RandomState(3)
    n = 40000
    ang = rng.uniform(0, 2 * np.pi, n); spd = rng.uniform(50, 5000, n)
    vx = (spd * np.cos(ang)).astype(np.float32); vy = (spd * np.sin(ang)).astype(np.float32)
    engine.set_states(pos, types, 1 << 1)
    a = engine.evaluate(0, vx, vy, W.GRID_SEARCH_COST, 0, final_pos=False)
    engine.level_select(512)
    beam_a = engine.fetch_beam()
    b = engine.evaluate(0, vx, vy, W.GRID_SEARCH_COST, 0, final_pos=False, select_k=512)
    beam_b = engine.fetch_beam()
    for key in ("cost", "flags", "potted", "n_events", "final_time"):
        assert np.array_equal(a[key].view(np.uint32) if a[key].dtype == np.float32 else a[key],
                              b[key].view(np.uint32) if b[key].dtype == np.float32 else b[key]), key
    assert len(beam_a["cost"]) == len(beam_b["cost"]) > 0
    for key in ("cost", "id", "potted"):
        assert np.array_equal(beam_a[key], beam_b[key]), key
    assert np.array_equal(beam_a["final_pos"].view(np.uint32), beam_b["final_pos"].view(np.uint32))


def test_two_engines_on_two_threads_do_not_share_state(table, port):
    """include/billiard_cuda.h: "all state lives in the engine ... different engines may be used from different threads at the
    same time".  Two engines with different ball counts, launch shapes and prune modes run levels concurrently from two
    threads (ctypes releases the GIL inside the calls); each must return what it returns alone."""
    import threading
    from billiard_ai_b200 import Engine, workloads as W
    from billiard_ai_b200.workloads import SNOOKER_SCORES

    def make(n_balls, seed, prune):
        e = Engine.from_table(table, SNOOKER_SCORES, cue_type=TYPE_CODE["WHITE"])
        e.set_prune(prune)
        pos, types = W.random_layout(n_balls, seed)
        return e, pos, types

    jobs = [make(16, 41, 0), make(9, 42, 1)]
    grids = [W.grid(96, 16), W.grid(64, 24)]

    def level(i):
        e, pos, types = jobs[i]
        e.set_states(pos, types, 1 << TYPE_CODE["RED"])
        e.evaluate_grid(*grids[i], W.GRID_SEARCH_COST)
        r = e.fetch()
        tc, ti = e.topk(64)
        return r["cost"].copy(), r["flags_raw"].copy(), r["final_pos"].copy(), tc, ti

    try:
        alone = [level(0), level(1)]
        results = [[None] * 6, [None] * 6]

        def worker(i):
            for k in range(6):
                results[i][k] = level(i)
        threads = [threading.Thread(target=worker, args=(i,)) for i in range(2)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        for i in range(2):
            for k in range(6):
                assert all(np.array_equal(a, b) for a, b in zip(alone[i], results[i][k])), (i, k)
        # and against the oracle, so that "the same" is also "right"
        e, pos, types = jobs[0]
        vx, vy = W.grid_velocities(*grids[0])
        ora = port.evaluate(pos[None], types[None], 1 << TYPE_CODE["RED"], 0, vx, vy, W.GRID_SEARCH_COST, 0, threads=4)
        clean = ora["min_gap"] >= 1e-9
        assert np.array_equal(alone[0][0][clean], ora["cost"][clean])
    finally:
        for e, _, _ in jobs:
            e.close()


@pytest.mark.parametrize("layout", [1, 2], ids=["fused-step-kernel", "merged-round-kernel"])
def test_experimental_round_layouts_are_bit_identical(engine, table, layout):
    """BCU_LAYOUT = 1 (bcu_step.cuh: a group of lanes owns a strike) and 2 (bcu_round.cuh: apply folded into items, double-buffered
    layers): kept as measured experiments (both slower).  They must return exactly what the default three-kernel
    round returns -- every output incl. the event log, both prune modes."""
    from billiard_ai_b200 import Engine, workloads as W
    pos, types = W.snooker22_open(31)
    dirs, speeds = W.grid(256, 24)
    vx, vy = W.grid_velocities(dirs, speeds)
    old = os.environ.get("BCU_LAYOUT")
    os.environ["BCU_LAYOUT"] = str(layout)
    try:
        other = Engine.from_table(table, SNOOKER_SCORES, cue_type=TYPE_CODE["WHITE"])
    finally:
        if old is None:
            os.environ.pop("BCU_LAYOUT", None)
        else:
            os.environ["BCU_LAYOUT"] = old
    try:
        for prune in (0, 1):
            res = []
            for e in (engine, other):
                e.set_prune(prune)
                e.set_states(pos, types, 1 << TYPE_CODE["RED"])
                res.append(e.evaluate(0, vx, vy, 100000, 7, final_pos=True, event_log=True))
            for key in ("cost", "flags_raw", "potted", "event_log"):
                assert np.array_equal(res[0][key], res[1][key]), (layout, prune, key)
            assert np.array_equal(res[0]["final_time"].view(np.uint32), res[1]["final_time"].view(np.uint32))
            assert np.array_equal(res[0]["final_pos"].view(np.uint32), res[1]["final_pos"].view(np.uint32))
    finally:
        engine.set_prune(0)
        other.close()


@pytest.mark.parametrize("knobs", [{"BCU_DENSE_MIN": "0"}, {"BCU_DENSE_MIN": "0", "BCU_JOBS_PER_STRIKE": "1"},
                                   {"BCU_DENSE_MIN_ITEMS": "0"}, {"BCU_DENSE_MIN_QUARTIC": "0"}],
                         ids=["both", "both-queue-overflow", "items-only", "quartic-only"])
def test_large_part_kernels_are_bit_identical(engine, table, knobs):
    """launch_pipeline takes other kernels for large parts: items_kernel_dense (compiled for more resident CTAs per SM) and
    quartic_kernel_wide (CTAs of 512 threads whose warps start every tile of jobs together; tiles of one job class).  The
    thresholds are engine settings read from the environment at bcu_create: an engine that uses those kernels on a SMALL launch
    must return exactly what the default engine returns -- every output incl. the event log, both prune modes -- also when the
    two-ended job queue overflows and the launch is repeated with a larger one."""
    from billiard_ai_b200 import Engine, workloads as W
    pos, types = W.snooker22_open(33)
    dirs, speeds = W.grid(192, 24)
    vx, vy = W.grid_velocities(dirs, speeds)
    old = {k: os.environ.get(k) for k in knobs}
    os.environ.update(knobs)
    try:
        other = Engine.from_table(table, SNOOKER_SCORES, cue_type=TYPE_CODE["WHITE"])
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    try:
        for prune in (0, 1):
            res = []
            for e in (engine, other):
                e.set_prune(prune)
                e.set_states(pos, types, 1 << TYPE_CODE["RED"])
                res.append(e.evaluate(0, vx, vy, 100000, 7, final_pos=True, event_log=True))
            for key in ("cost", "flags_raw", "potted", "event_log"):
                assert np.array_equal(res[0][key], res[1][key]), (knobs, prune, key)
            assert np.array_equal(res[0]["final_time"].view(np.uint32), res[1]["final_time"].view(np.uint32))
            assert np.array_equal(res[0]["final_pos"].view(np.uint32), res[1]["final_pos"].view(np.uint32))
    finally:
        engine.set_prune(0)
        other.close()
