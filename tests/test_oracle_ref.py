"""Pins the reference oracle (oracle/_ref/libref_oracle.so = the reference's own sources) against every
golden vector the reference holds for this path."""
import json
import os

import numpy as np
import pytest

from oracle_lib import GOLDEN, TYPE_CODE


def _close(z, re, im, tol=1e-4):
    # tolerance and comparison of test/billiard_physics/src/billiard_physics_test.cpp:165-171
    return abs(z.real - re) < tol and abs(abs(z.imag) - abs(im)) < tol


@pytest.mark.parametrize("oracle", ["ref", "port"])
def test_quartic_known_answers(oracle, request):
    o = request.getfixturevalue(oracle)
    # billiard_physics_test.cpp:84-143
    r = o.quartic(1, -4, 0, 8, 1)
    for z, e in zip(r, [-1.1753, -0.12603, 2.1260, 3.1753]):
        assert _close(z, e, 0.0)
    r = o.quartic(1, -4, 3, 8, 1)
    assert _close(r[0], -0.91904, 0) and _close(r[1], -0.13283, 0)
    assert _close(r[2], 2.5259, 1.3459) and _close(r[3], 2.5259, 1.3459)
    r = o.quartic(0.009604, -0.0274841, 0.0327719, -0.018757, 0.00174089)
    assert _close(r[0], 0.11314, 0) and _close(r[3], 1.31773, 0)
    assert _close(r[1], 0.715435, 0.839049) and _close(r[2], 0.715435, 0.839049)


def test_time_to_collision_known_answer(ref):
    # billiard_physics_test.cpp:145-163 (7-argument form of the current header), expects 0.11314 +- 1e-4
    t = ref.time_to_collision((94.7511, -171.576), (-67.7881, 122.751), (-868.681, 365.701), (-0.0, -0.0), (0, 0),
                              (-899.492, 425.088), 52.3)
    assert t is not None and abs(t - 0.11314) < 1e-4


def _shots():
    with open(os.path.join(GOLDEN, "filmed_shots.json")) as f:
        return json.load(f)["tests"]


@pytest.mark.parametrize("oracle", ["ref", "port"])
@pytest.mark.parametrize("shot", [s for s in _shots() if s["enabled"]], ids=lambda s: s["name"])
def test_filmed_shot_structure(oracle, shot, request):
    """test/billiard_search/src/simulation_vs_reality.cpp: asserted layer counts and node kinds."""
    o = request.getfixturevalue(oracle)
    pos = np.array([b["position"] for b in shot["balls"]], dtype=np.float32)
    types = np.array([TYPE_CODE[b["type"]] for b in shot["balls"]], dtype=np.int32)
    out = o.simulate_trace(pos, types, *shot["velocity"])
    nl, rec = out[0], out[2]
    if shot["expectedLayers"]:
        assert nl in shot["expectedLayers"]
    var_index = {b["var"]: i for i, b in enumerate(shot["balls"])}
    for e in shot["expects"]:
        if e.get("onlyIfLayers", nl) != nl or e.get("notIfLayers", -1) == nl or e["layer"] >= nl:
            continue  # expectations inside `if (layers == K) ... else ...` of the reference test
        r = rec[e["layer"], var_index[e["ball"]]]
        assert int(r[0]) == e["kind"], (shot["name"], e)
        if "afterRolling" in e:
            assert bool(r[5]) == e["afterRolling"]


def test_thesis_cost_lists(ref):
    """doc_source/.../04_suche.tex: the published ranked totals are reproduced exactly."""
    with open(os.path.join(GOLDEN, "thesis_costs.json")) as f:
        thesis = json.load(f)
    with open(os.path.join(GOLDEN, "layouts.json")) as f:
        lay = json.load(f)["situation1"]
    pos = np.array([b["position"] for b in lay["balls"]], dtype=np.float32)
    types = np.array([TYPE_CODE[b["type"]] for b in lay["balls"]], dtype=np.int32)
    none = np.uint64(2**64 - 1)
    ref.set_toggles(0, 0)
    pl = ref.plan(pos, types, 1 << TYPE_CODE["RED"])
    got = sorted(int(c) for c in pl["result_cost"] if c != none)
    assert got[:10] == thesis["lists"]["s1_with_banks"]
    assert set(thesis["lists"]["s1_single_no_banks"]) <= set(got)
    ref.set_toggles(0, 1)
    pl = ref.plan(pos, types, 1 << TYPE_CODE["RED"])
    got = sorted(int(c) for c in pl["result_cost"] if c != none)
    assert got[:10] == thesis["lists"]["s1_three_speeds"]
    # every logged simulation cost appears as (result - base) of some candidate
    sim_costs = {int(r) - int(b) for r, b in zip(pl["result_cost"], pl["base_cost"]) if r != none}
    for log in thesis["logs"]:
        if 400 < log["line"] < 620:
            assert log["simulation"][0][0] in sim_costs


def test_thesis_two_break_list(ref):
    """04_suche.tex:1114-1246: the two-break totals (183281, 185129, ...) come out of the reference's own search -- expand()
    run best-first to exhaustion with depth search on, BREAKS = 2 -- with the reference's own snooker rules between the breaks
    (nextSearch / stateAfterBreak, oracle/_ref/libref_rules.so).  Pins the multi-break path of the oracle."""
    import ctypes as C
    from oracle_lib import ORACLE_DIR, TYPE_NAMES
    so = os.path.join(ORACLE_DIR, "_ref", "libmirror_check.so")
    if not os.path.exists(so):
        pytest.skip("libmirror_check.so not built (needs /root/reference at build time)")
    from billiard_ai_b200.rules import SnookerRules
    from test_rules import RefRules
    RefRules(SnookerRules())
    lib = C.CDLL(so)
    lib.mirror_reference_search.restype = C.c_int64
    with open(os.path.join(GOLDEN, "thesis_costs.json")) as f:
        want = json.load(f)["lists"]["s2_two_breaks"]
    with open(os.path.join(GOLDEN, "layouts.json")) as f:
        lay = json.load(f)["situation2"]
    pos = np.array([b["position"] for b in lay["balls"]], dtype=np.float32)
    types = np.array([TYPE_CODE[b["type"]] for b in lay["balls"]], dtype=np.int32)
    searched = np.array([TYPE_CODE["RED"]], dtype=np.int32)
    names = (C.c_char_p * len(TYPE_NAMES))(*[t.encode() for t in TYPE_NAMES])
    costs = np.zeros(1024, np.uint64)
    breaks = np.zeros(1024, np.int32)
    p = lambda a, t: a.ctypes.data_as(C.POINTER(t))   # noqa: E731
    n = lib.mirror_reference_search(C.c_int(len(types)), p(pos, C.c_float), names, C.c_int(len(TYPE_NAMES)), p(types, C.c_int32),
                                    C.c_int(1), p(searched, C.c_int32), C.c_int(1), C.c_int(0), C.c_int(2), C.c_int64(1024),
                                    p(costs, C.c_uint64), p(breaks, C.c_int32))
    assert n >= len(want)
    assert costs[:len(want)].tolist() == want and (breaks[:n] == 2).all()
