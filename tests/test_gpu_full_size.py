"""Full-size parity: the configurations bench.py times, compared strike by strike with the REFERENCE ITSELF
(oracle/_ref/libref_oracle.so = the reference's unmodified sources, all host threads), not through properties.

  * BASELINE config 2 (the benched workload): all 65 536 strikes of the 22-ball layout x (1024 directions x 64 speeds) --
    flags, event counts, potted masks, costs, bit-equal end times and end positions, the identical top-1024 (cost, id)
    list and the identical best strike.  Both with the reach bound off (the C ABI's default) and on (what bench.py's
    `value` runs with).
  * the 16-ball level the north_star target is quoted on, and a slice of config 5 (random 16-ball tables x 4 096 strikes).
  * the multi-GPU merge kernel on one GPU (bcu_test_merge_records) against the numpy statement of the merge rule.

Near ties (two candidate events within 1e-9 s, north_star's tie clause) are masked with the port oracle's `min_gap`
(the port is bit-identical to the reference build, tests/test_oracle_port.py, and carries the tie information the
reference build has no output for); everything else must agree 100 %."""
import os

import numpy as np
import pytest

from oracle_lib import TYPE_CODE, SNOOKER_SCORES, F_VALID
from parity import TIE_EPS

pytestmark = pytest.mark.gpu
THREADS = os.cpu_count() or 4
NONE = np.uint64(2**64 - 1)


@pytest.fixture(scope="module")
def engine(table):
    from billiard_ai_b200 import Engine
    e = Engine.from_table(table, SNOOKER_SCORES, cue_type=TYPE_CODE["WHITE"])
    yield e
    e.close()


def assert_identical(got, want, clean, what):
    """Every output of every tie-free strike: integers equal, floats bit-equal."""
    for key in ("flags", "n_events", "potted", "cost"):
        bad = np.nonzero((got[key] != want[key]) & clean)[0]
        assert len(bad) == 0, (what, key, len(bad), bad[:8].tolist())
    bad = np.nonzero((got["final_time"].view(np.uint32) != want["final_time"].view(np.uint32)) & clean)[0]
    assert len(bad) == 0, (what, "final_time", len(bad), bad[:8].tolist())
    n = len(clean)
    diff = (got["final_pos"].view(np.uint32) != want["final_pos"].view(np.uint32)).reshape(n, -1).any(axis=1)
    bad = np.nonzero(diff & clean)[0]
    assert len(bad) == 0, (what, "final_pos", len(bad), bad[:8].tolist())


def reference_topk(res, k, clean=None):
    """Ascending (cost, id) over the valid strikes: the reference's solution order (process.hpp:19-25) + id tie-break."""
    ids = np.nonzero((res["flags"] & F_VALID) != 0)[0]
    order = ids[np.lexsort((ids, res["cost"][ids]))][:k]
    return order.astype(np.uint64), res["cost"][order]


@pytest.mark.parametrize("prune", [0, 1])
def test_config2_every_strike_against_the_reference_build(engine, ref, port, prune):
    from billiard_ai_b200 import workloads as W
    pos, types = W.snooker22_open()
    dirs, speeds = W.grid(1024, 64)
    vx, vy = W.grid_velocities(dirs, speeds)
    mask = 1 << TYPE_CODE["RED"]
    want = ref.evaluate(pos[None], types[None], mask, 0, vx, vy, W.GRID_SEARCH_COST, 0, threads=THREADS)
    tie = port.evaluate(pos[None], types[None], mask, 0, vx, vy, W.GRID_SEARCH_COST, 0, threads=THREADS)
    for key in ("flags", "n_events", "potted", "cost"):   # the port is the reference, bit for bit, on this workload too
        assert np.array_equal(tie[key], want[key]), key
    clean = tie["min_gap"] >= TIE_EPS
    assert clean.sum() >= 0.99 * len(vx)
    engine.set_prune(prune)
    try:
        engine.set_states(pos, types, mask)
        assert engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST) == 65536
        got = engine.fetch()
        assert_identical(got, want, clean, f"config 2, prune {prune}")
        # the selection: identical top-1024 list and best strike (a near-tie strike may only differ if it is in neither list)
        engine.level_select(1024)
        beam = engine.fetch_beam()
        ids, costs = reference_topk(want, 1024)
        assert len(beam["id"]) == len(ids) == 1024
        assert clean[ids.astype(np.int64)].all(), "a selected strike is a near tie: compare up to it"
        assert np.array_equal(beam["id"], ids) and np.array_equal(beam["cost"], costs)
        assert int(beam["id"][0]) == int(ids[0]) and int(beam["cost"][0]) == int(costs[0])       # the best strike
        sel = ids.astype(np.int64)
        assert np.array_equal(beam["final_pos"].view(np.uint32), want["final_pos"][sel].view(np.uint32))
        assert np.array_equal(beam["potted"], want["potted"][sel])
        print(f"config 2, prune {prune}: 65536 strikes, {int(clean.sum())} tie-free, all identical; best strike id "
              f"{int(ids[0])} cost {int(costs[0])}")
    finally:
        engine.set_prune(0)


@pytest.mark.parametrize("prune", [0, 1])
def test_sixteen_ball_level_every_strike_against_the_reference_build(engine, ref, port, prune):
    """The layout the north_star throughput target is quoted on: 16 balls (seed 16), 1024 x 64 strikes."""
    from billiard_ai_b200 import workloads as W
    pos, types = W.random_layout(16, 16)
    dirs, speeds = W.grid(1024, 64)
    vx, vy = W.grid_velocities(dirs, speeds)
    want = ref.evaluate(pos[None], types[None], 2, 0, vx, vy, W.GRID_SEARCH_COST, 0, threads=THREADS)
    tie = port.evaluate(pos[None], types[None], 2, 0, vx, vy, W.GRID_SEARCH_COST, 0, threads=THREADS)
    clean = tie["min_gap"] >= TIE_EPS
    engine.set_prune(prune)
    try:
        engine.set_states(pos, types, 2)
        engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
        got = engine.fetch()
        assert_identical(got, want, clean, f"16 balls, prune {prune}")
        engine.level_select(1024)
        beam = engine.fetch_beam()
        ids, costs = reference_topk(want, 1024)
        if clean[ids.astype(np.int64)].all():
            assert np.array_equal(beam["id"], ids) and np.array_equal(beam["cost"], costs)
    finally:
        engine.set_prune(0)


@pytest.mark.parametrize("prune", [0, 1])
def test_config5_slice_against_the_reference_build(engine, ref, port, prune):
    """BASELINE config 5 (10 000 random 16-ball tables x 4 096 strikes): tables 1000..1047 of the sweep, every strike."""
    from billiard_ai_b200 import workloads as W
    nt = 48
    pos, types = W.random_tables(nt, 16, seed0=1000)
    dirs, speeds = W.grid(64, 64)
    vx1, vy1 = W.grid_velocities(dirs, speeds)
    cs = np.repeat(np.arange(nt, dtype=np.int32), len(vx1))
    vx, vy = np.tile(vx1, nt), np.tile(vy1, nt)
    want = ref.evaluate(pos, types, 2, cs, vx, vy, W.GRID_SEARCH_COST, 0, threads=THREADS)
    tie = port.evaluate(pos, types, 2, cs, vx, vy, W.GRID_SEARCH_COST, 0, threads=THREADS)
    clean = tie["min_gap"] >= TIE_EPS
    assert clean.sum() >= 0.98 * len(vx)
    engine.set_prune(prune)
    try:
        engine.set_states(pos, types, 2)
        assert engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST) == nt * 4096
        got = engine.fetch()
        assert_identical(got, want, clean, f"config 5 slice, prune {prune}")
    finally:
        engine.set_prune(0)


# ---- the multi-GPU merge on one GPU ---------------------------------------------------------------------------------
def _synthetic_blocks(rng, G, k, nb, fill):
    """G sorted record blocks as G ranks would pack them: many equal costs (the id decides), ranks with fewer than k
    valid entries and empty ranks (`fill[r]` = number of valid entries of rank r)."""
    import merge_restatement as D
    blocks = []
    next_id = 0
    for r in range(G):
        m = fill[r]
        span = max(m, 1) * 3
        ids = np.sort(rng.choice(span, size=m, replace=False)).astype(np.uint64) + np.uint64(next_id)
        next_id += span                      # contiguous id ranges per rank, as shard_range gives them
        cost = rng.randint(0, 40, m).astype(np.uint64) * 500
        order = np.lexsort((ids, cost))
        blk = np.zeros(D.block_bytes(k, nb), np.uint8)
        v = D.view_block(blk, k, nb)
        v["cost"][:] = D.NONE; v["id"][:] = D.NONE
        v["cost"][:m] = cost[order]; v["id"][:m] = ids[order]
        v["potted"][:m] = rng.randint(0, 1 << nb, m).astype(np.uint32)
        v["time"][:m] = rng.rand(m).astype(np.float32)
        v["pos"][:m] = rng.rand(m, nb, 2).astype(np.float32)
        blocks.append(blk)
    return blocks


@pytest.mark.parametrize("G,k,nb", [(2, 1024, 22), (4, 1024, 22), (8, 1024, 22), (8, 100, 16), (3, 7, 5), (1, 64, 22)])
def test_merge_records_kernel_against_the_numpy_merge(engine, G, k, nb):
    """merge_records_kernel (csrc/bcu_merge.cu) only runs behind ncclAllGather with more than one rank; the test hook
    feeds it G host blocks.  Must equal merge_restatement.merge_blocks byte for byte: full ranks, short ranks, empty ranks,
    fewer than k valid entries in total."""
    import merge_restatement as D
    rng = np.random.RandomState(G * 1000 + k)
    fills = [[k] * G,                                              # every rank full
             [max(0, k - 17 * r) for r in range(G)],               # short ranks
             [0] + [k // 3] * (G - 1),                             # an empty rank; fewer than k in total when G is small
             [0] * G,                                              # nothing valid anywhere
             [1] * G]
    for fill in fills:
        blocks = _synthetic_blocks(rng, G, k, nb, fill)
        got = engine.test_merge_records(blocks, k, nb)
        want = D.merge_blocks(blocks, k, nb)
        g, w = D.view_block(got, k, nb), D.view_block(want, k, nb)
        for key in ("cost", "id", "potted"):
            assert np.array_equal(g[key], w[key]), (fill, key)
        assert np.array_equal(g["time"].view(np.uint32), w["time"].view(np.uint32)), fill
        assert np.array_equal(g["pos"].view(np.uint32), w["pos"].view(np.uint32)), fill


def test_trace_does_not_disturb_the_uploaded_states(engine):
    """bcu_simulate_trace keeps its one-state table in buffers of its own: a trace between
    bcu_set_states and bcu_evaluate_grid must not change what the grid is evaluated on."""
    from billiard_ai_b200 import workloads as W
    pos, types = W.random_layout(16, 16)
    dirs, speeds = W.grid(64, 8)
    engine.set_states(pos, types, 2)
    engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
    a = engine.fetch()
    engine.set_states(pos, types, 2)
    other, otypes = W.random_layout(5, 99)
    nl, _, _ = engine.simulate_trace(other, otypes, 1500.0, 300.0)
    assert nl >= 2
    engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
    b = engine.fetch()
    for key in ("cost", "flags", "potted", "final_pos"):
        assert np.array_equal(a[key], b[key]), key


def test_empty_shard_and_grid_limits(engine):
    """A rank whose shard is empty (more ranks than candidates) still takes part in the level: zero candidates, an empty
    local beam, a closed timer.  n_dirs * n_speeds beyond 2^31 - 1 is rejected (candidate decode is 32-bit per state)."""
    from billiard_ai_b200 import workloads as W, BcuError
    pos, types = W.random_layout(6, 3)
    dirs, speeds = W.grid(8, 4)
    engine.set_states(pos, types, 2)
    assert engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST, first=32, count=0) == 0
    assert engine.last_kernel_ms() >= 0.0
    engine.level_select(16)
    assert len(engine.fetch_beam()["cost"]) == 0
    import ctypes as C
    rc = engine.lib.bcu_evaluate_grid(engine.h, C.c_int32(1 << 20), C.c_int32(1 << 12), dirs.ctypes.data_as(C.POINTER(C.c_float)),
                                      speeds.ctypes.data_as(C.POINTER(C.c_float)), C.c_uint64(1), C.c_int64(0), C.c_int64(1))
    assert rc == 2   # BCU_ERR_INVALID
