"""Regression vectors of the reach bound (csrc/bcu_items.cuh; DESIGN.md section 4, "audited, not proven").

tests/golden/prune_regressions.json holds the strikes on which a weakened bound is known to change the result -- the one
failure of the round-1 audits and the cases found with the guard-variant builds (tools/find_guard_cases.py) -- together with
the reference's own results for them (oracle/_ref, generated in the build container).  On every vector

  * prune modes 0 and 1 of the product reproduce the reference, event log included, and the audit mode reports no failure;
  * the x 10 and / 10 builds of every threshold still do (the margin of the thresholds is at least a decade on the vectors);
  * the build with the guard REMOVED that the vector names in `defeats` does NOT (so the vector demonstrably exercises it)."""
import json
import os

import numpy as np
import pytest

from oracle_lib import GOLDEN, TYPE_CODE, SNOOKER_SCORES

pytestmark = pytest.mark.gpu


def _cases():
    with open(os.path.join(GOLDEN, "prune_regressions.json")) as f:
        return json.load(f)["cases"]


def _inputs(c):
    f32 = lambda key: np.array(c[key], np.uint32).view(np.float32)   # noqa: E731
    return f32("pos_bits").reshape(-1, 2), np.array(c["types"], np.int32), f32("vx_bits"), f32("vy_bits")


def _matches_reference(out, c):
    r = c["reference"]
    same = (np.array_equal(out["cost"], np.array(r["cost"], np.uint64)) and np.array_equal(out["flags"], np.array(r["flags"], np.uint32))
            and np.array_equal(out["n_events"], np.array(r["n_events"], np.int32))
            and np.array_equal(out["potted"], np.array(r["potted"], np.uint32))
            and np.array_equal(out["final_time"].view(np.uint32), np.array(r["final_time_bits"], np.uint32)))
    got, want = out["event_log"], np.array(r["event_log"], np.int32)
    coll = want[..., 0] == 0   # the reference's {ball1, ball2} order depends on hash-map iteration: compare as a set
    lo = np.where(coll, np.minimum(want[..., 1], want[..., 2]), want[..., 1])
    hi = np.where(coll, np.maximum(want[..., 1], want[..., 2]), want[..., 2])
    return same and np.array_equal(got[..., 0], want[..., 0]) and np.array_equal(got[..., 1], lo) and \
        np.array_equal(got[..., 2], hi) and np.array_equal(got[..., 3], want[..., 3])


def _engine(table, lib_path=None):
    from billiard_ai_b200 import Engine
    return Engine.from_table(table, SNOOKER_SCORES, cue_type=TYPE_CODE["WHITE"], lib_path=lib_path)


def _run(e, c, mode):
    pos, types, vx, vy = _inputs(c)
    e.set_prune(mode)
    e.set_states(pos, types, 2)
    return e.evaluate(0, vx, vy, 100000, 0, final_pos=False, event_log=True)


@pytest.mark.parametrize("case", _cases(), ids=lambda c: c["name"])
def test_product_reproduces_the_reference_on_the_regression_vectors(table, case):
    e = _engine(table)
    try:
        e.enable_counters(True)
        for mode in (0, 1):
            assert _matches_reference(_run(e, case, mode), case), (case["name"], mode)
        _run(e, case, 2)
        assert e.prune_counters()["audit_failures"] == 0
    finally:
        e.close()


def test_threshold_sweep_and_removed_guards(table):
    """Thresholds x 10 and / 10 keep every vector green; removing the guard a vector names breaks it."""
    from billiard_ai_b200.build import GUARD_VARIANTS, variant_path
    cases = _cases()
    missing = [n for n in GUARD_VARIANTS if not os.path.exists(variant_path(n))]
    if missing:
        pytest.skip(f"guard-variant builds not present: {missing} (python -m billiard_ai_b200.build --variants)")
    report = {}
    for name in GUARD_VARIANTS:
        e = _engine(table, variant_path(name))
        try:
            report[name] = {c["name"]: bool(_matches_reference(_run(e, c, 1), c)) for c in cases}
            for c in cases:   # mode 0 of any build is the reference
                assert _matches_reference(_run(e, c, 0), c), (name, c["name"])
        finally:
            e.close()
    print(json.dumps(report, indent=1))
    for name in GUARD_VARIANTS:
        for c in cases:
            if name in c["defeats"]:
                assert not report[name][c["name"]], f"{c['name']} no longer exercises the guard removed in build {name}"
            elif name.endswith(("_x10", "_d10")):
                assert report[name][c["name"]], f"threshold variant {name} changes the result of {c['name']}"
