"""The search() drop-in of the C++ host mirror (billiard_ai_b200/host/billiard_cuda_search.hpp: cuda::search / SearchFront)
against the reference's own search run to exhaustion (tests/host_mirror_check.cpp, compiled against the reference's headers
into oracle/_ref/libmirror_check.so and linked with the reference build, the reference's rules and the product).

Reference side: a deterministic best-first loop over the reference's expand() -- its planner and its CPU simulate() -- until no
node is left (the reference's own search() is timing dependent: worker threads, 100 ms sync, 10 s budget).  Drop-in side: the
reference's planner on the host, every SIMULATION node evaluated on the device, the Configuration's own std::function rules
between breaks, results as std::vector<std::vector<node::System>>.  Compared: solution costs in order, and every returned
solution's Systems (one per break) node by node and bit by bit with a reference solution of the same cost.

Layouts: test/billiard_search `00_test_01`, the thesis' Situation 1 and Situation 2 (doc_source/.../04_suche.tex), whose
published totals (72206, 76365, ...; two breaks: 183281, ...) are the known answers (tests/golden/thesis_costs.json)."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from oracle_lib import GOLDEN, ORACLE_DIR, TYPE_CODE, TYPE_NAMES

pytestmark = pytest.mark.gpu
SO = os.path.join(ORACLE_DIR, "_ref", "libmirror_check.so")


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


@pytest.fixture(scope="module")
def mirror(ref):
    """`ref` has configured the reference build (table, rules) that libmirror_check.so links against."""
    if not os.path.exists(SO):
        pytest.skip("libmirror_check.so not built (needs /root/reference at build time)")
    from billiard_ai_b200.rules import SnookerRules
    from test_rules import RefRules
    RefRules(SnookerRules())   # refrules_configure: the searchConfig of the reference's snooker rules (same process, same library)
    return C.CDLL(SO)


def _layout(name):
    with open(os.path.join(GOLDEN, "layouts.json")) as f:
        lay = json.load(f)[name]
    pos = np.array([b["position"] for b in lay["balls"]], dtype=np.float32)
    types = np.array([TYPE_CODE[b["type"]] for b in lay["balls"]], dtype=np.int32)
    searched = np.array([TYPE_CODE[t] for t in lay.get("searchTypes", ["RED"])], dtype=np.int32)
    return pos, types, searched


def _search(lib, name, solutions, depth, vm, rules_kind, max_costs=64, batch=0):
    pos, types, searched = _layout(name)
    names = (C.c_char_p * len(TYPE_NAMES))(*[t.encode() for t in TYPE_NAMES])
    out = np.zeros(10, np.int64)
    costs = np.zeros((2, max_costs), np.uint64)
    ms = np.zeros(2, np.float64)
    rc = lib.mirror_search_check(C.c_int(len(types)), _p(pos, C.c_float), names, C.c_int(len(TYPE_NAMES)), _p(types, C.c_int32),
                                 C.c_int(len(searched)), _p(searched, C.c_int32), C.c_int(solutions), C.c_int(depth), C.c_int(vm),
                                 C.c_int(rules_kind), C.c_int(batch), _p(out, C.c_int64), C.c_int(max_costs), _p(costs, C.c_uint64),
                                 _p(ms, C.c_double))
    assert rc == 0
    keys = ("ref_solutions", "solutions", "cost_mismatches", "system_mismatches", "ref_expansions", "expansions", "simulations",
            "batches", "enough_solutions", "no_work")
    r = dict(zip(keys, out.tolist()))
    none = np.uint64(2**64 - 1)
    r["ref_costs"] = [int(c) for c in costs[0] if c != none]
    r["costs"] = [int(c) for c in costs[1] if c != none]
    r["ms_reference"], r["ms_dropin"] = ms.tolist()
    print(name, {k: v for k, v in r.items() if k not in ("ref_costs", "costs")}, r["costs"][:8])
    return r


def _thesis():
    with open(os.path.join(GOLDEN, "thesis_costs.json")) as f:
        return json.load(f)["lists"]


@pytest.mark.parametrize("name", ["test_01", "situation1", "situation2"])
@pytest.mark.parametrize("vm", [0, 1])
def test_single_break_search_returns_the_reference_solutions(mirror, name, vm):
    """Config 1: every solution the reference's search finds, ascending, with identical Systems; exhaustive request."""
    r = _search(mirror, name, 65535, 0, vm, 0)
    assert r["ref_solutions"] > 0 and r["solutions"] == r["ref_solutions"]
    assert r["cost_mismatches"] == 0 and r["system_mismatches"] == 0
    assert r["costs"] == r["ref_costs"][:len(r["costs"])] and r["no_work"] == 1
    if name == "situation1":
        want = _thesis()["s1_three_speeds" if vm else "s1_with_banks"]
        assert r["costs"][:len(want)] == want   # 72206, 76365, 79920, ... as printed in the thesis


def test_requesting_few_solutions_ends_early_with_the_cheapest(mirror):
    """`solutions` = 3 (process.hpp:188-198 enoughSolutions): the three cheapest solutions that exist, found with fewer
    expansions than the exhaustive search needs."""
    full = _search(mirror, "situation1", 65535, 0, 1, 0)
    for batch in (0, 8, 1):   # default (32), and launches of 8 strikes / 1 strike (the reference's own granularity)
        few = _search(mirror, "situation1", 3, 0, 1, 0, batch=batch)
        assert few["cost_mismatches"] == 0 and few["system_mismatches"] == 0
        assert few["solutions"] >= 3 and few["costs"][:3] == full["costs"][:3]
        assert few["enough_solutions"] == 1 and few["expansions"] < full["expansions"] and few["simulations"] < full["simulations"]


@pytest.mark.parametrize("rules_kind", [1, 2], ids=["repeat", "snooker"])
def test_two_break_search_returns_the_reference_solutions(mirror, rules_kind):
    """Depth search (BREAKS = 2, billiard_search.cpp:15, :1407-1462) with the rules between breaks as the Configuration's own
    std::functions: a 'repeat' rule set and the reference's snooker rules (nextSearch / stateAfterBreak)."""
    r = _search(mirror, "situation2", 10, 1, 0, rules_kind)
    assert r["ref_solutions"] > 0 and r["solutions"] >= min(10, r["ref_solutions"])
    assert r["cost_mismatches"] == 0 and r["system_mismatches"] == 0
    assert r["costs"][:10] == r["ref_costs"][:10]
    if rules_kind == 2:
        want = _thesis()["s2_two_breaks"]
        assert r["costs"][:len(want)] == want, (want, r["costs"][:8])   # 183281, 185129, ... as printed in the thesis


def test_front_replaces_requests_and_honours_budget(mirror):
    """test/process/src/process_test.cpp:29-132 in spirit: a newer request replaces the one in flight (its future breaks its
    promise, exactly as the reference's dropped std::promise does), an invalid search and `solutions` = 0 return at once,
    a 20 ms budget ends an exhaustive request early, and the free function search() answers."""
    pos, types, searched = _layout("situation2")
    names = (C.c_char_p * len(TYPE_NAMES))(*[t.encode() for t in TYPE_NAMES])
    out = np.zeros(8, np.int64)
    rc = mirror.mirror_front_check(C.c_int(len(types)), _p(pos, C.c_float), names, C.c_int(len(TYPE_NAMES)), _p(types, C.c_int32),
                                   C.c_int(len(searched)), _p(searched, C.c_int32), _p(out, C.c_int64))
    assert rc == 0
    print(out.tolist())
    assert out[0] == 1, "the replaced request's future did not break its promise"
    assert out[1] >= 1 and out[2] == 1 and out[6] == 1
    assert out[3] == 1 and out[4] < 5000
    assert out[5] >= 3
    assert out[7] == 1, "searchOnly() differs from search()"


@pytest.mark.parametrize("case", [("test_01", 0, 1, 0), ("situation1", 0, 0, 0), ("situation1", 0, 1, 0), ("situation2", 1, 0, 2), ("situation2", 1, 0, 1)],
                         ids=["test_01-3speeds", "situation1", "situation1-3speeds", "situation2-two-breaks-snooker", "situation2-two-breaks-repeat"])
def test_search_with_the_device_planner_returns_the_reference_solutions(mirror, case):
    """SearchOptions::devicePlanner: nothing of the reference's planner runs on the host -- per break every state is planned
    (bcu_plan) and simulated (bcu_evaluate_plan) on the device, the rules between breaks stay the Configuration's own
    std::functions.  Exhaustive per break, so it must return ALL solutions of the reference's search run to exhaustion, in
    cost order, with identical Systems."""
    name, depth, vm, rules_kind = case
    r = _search(mirror, name, 65535, depth, vm, rules_kind, batch=-1)
    assert r["ref_solutions"] > 0 and r["solutions"] == r["ref_solutions"]
    assert r["cost_mismatches"] == 0 and r["system_mismatches"] == 0
    assert r["costs"] == r["ref_costs"][:len(r["costs"])]
