"""The reference's geometric planner restated (billiard_ai_b200/csrc/bcu_planner.cuh).

CPU tier (no device): the header compiled for the host inside oracle/_ref/libplanner_check.so (tests/planner_check.cpp)
against the reference's own expand() run over every SEARCH node: the two multisets of strikes -- BALL_SHOT velocity bits,
_searchCost, accumulated _cost of every SIMULATION node -- must be identical, on the reference's own test layout, the thesis'
two situations and seeded random tables of 5 to 22 balls, with and without the faster-strike variants, at break 1 (cost 0)
and as a later break (accumulated cost).

GPU tier: plan_kernel (bcu_plan.cu) through the C ABI -- bcu_set_planner / bcu_plan / bcu_fetch_plan / bcu_evaluate_plan -- on
a beam of states at once: the plan of every state equals the reference's, and simulating it gives the reference's solution
costs (the thesis' 72206, 76365, ...)."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from oracle_lib import GOLDEN, ORACLE_DIR, TYPE_CODE, TYPE_NAMES

SO = os.path.join(ORACLE_DIR, "_ref", "libplanner_check.so")
NONE = np.uint64(2**64 - 1)


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


@pytest.fixture(scope="module")
def check(ref):
    """`ref` has configured the reference build (table, rules) the check library links against."""
    if not os.path.exists(SO):
        pytest.skip("libplanner_check.so not built (needs /root/reference at build time)")
    return C.CDLL(SO)


def _layout(name):
    with open(os.path.join(GOLDEN, "layouts.json")) as f:
        lay = json.load(f)[name]
    pos = np.array([b["position"] for b in lay["balls"]], dtype=np.float32)
    types = np.array([TYPE_CODE[b["type"]] for b in lay["balls"]], dtype=np.int32)
    searched = np.array([TYPE_CODE[t] for t in lay.get("searchTypes", ["RED"])], dtype=np.int32)
    return pos, types, searched


def _reference_plan(lib, pos, types, searched, vm, base=0, max_out=4096):
    names = (C.c_char_p * len(TYPE_NAMES))(*[t.encode() for t in TYPE_NAMES])
    out = np.zeros(8, np.int64)
    vel = np.zeros((max_out, 2), np.float32)
    cost = np.zeros((max_out, 2), np.uint64)
    rc = lib.planner_check(C.c_int(len(types)), _p(pos, C.c_float), names, C.c_int(len(TYPE_NAMES)), _p(types, C.c_int32),
                           C.c_int(len(searched)), _p(searched, C.c_int32), C.c_int(vm), C.c_uint64(base), _p(out, C.c_int64),
                           C.c_int64(max_out), _p(vel, C.c_float), _p(cost, C.c_uint64))
    assert rc == 0
    n = int(out[0])
    assert n <= max_out
    return dict(n_ref=n, n_got=int(out[1]), only_ref=int(out[2]), only_got=int(out[3]), expansions=int(out[4]),
                vx=vel[:n, 0].copy(), vy=vel[:n, 1].copy(), search_cost=cost[:n, 0].copy(), cost=cost[:n, 1].copy())


@pytest.mark.parametrize("name", ["test_01", "situation1", "situation2"])
@pytest.mark.parametrize("vm", [0, 1])
def test_restatement_equals_the_reference_planner_on_its_layouts(check, name, vm):
    pos, types, searched = _layout(name)
    r = _reference_plan(check, pos, types, searched, vm)
    print(name, vm, {k: v for k, v in r.items() if not isinstance(v, np.ndarray)})
    assert r["n_ref"] > 0 and r["n_got"] == r["n_ref"] and r["only_ref"] == 0 and r["only_got"] == 0


def test_restatement_equals_the_reference_planner_on_random_tables(check):
    from billiard_ai_b200 import workloads as W
    total = 0
    for seed in range(30):
        n = [5, 8, 12, 16, 22][seed % 5]
        pos, types = W.random_layout(n, 100 + seed) if n < 22 else W.snooker22_open(100 + seed)
        searched = np.array([2, 3, 4, 5, 6, 7] if n == 22 and seed % 2 else [1], np.int32)
        r = _reference_plan(check, pos, types, searched, seed % 2, base=seed * 1000)
        assert r["n_got"] == r["n_ref"] and r["only_ref"] == 0 and r["only_got"] == 0, (seed, n, r)
        total += r["n_ref"]
    assert total > 1000


def test_golden_rail_orders_are_the_reference_containers_orders(check, table):
    """tests/golden/table.json railMapOrder / railSetOrder (inputs of the device planner) against the reference build here."""
    ids = [r["id"] for r in table["rails"]]
    rail_order = np.zeros(len(ids), np.int32)
    set_order = np.zeros(len(ids), np.int32)
    n_set = C.c_int32(0)
    n = check.planner_orders("\n".join(ids).encode(), _p(rail_order, C.c_int32), _p(set_order, C.c_int32), C.byref(n_set))
    assert n == len(ids)
    assert [ids[i] for i in rail_order] == table["railMapOrder"]
    assert [ids[i] for i in set_order[:n_set.value]] == table["railSetOrder"]


def _multiset(vx, vy, sc, cost):
    return sorted(zip(vx.view(np.uint32).tolist(), vy.view(np.uint32).tolist(), sc.tolist(), cost.tolist()))


@pytest.mark.gpu
def test_device_plan_of_a_beam_equals_the_reference(check, table, ref):
    """plan_kernel over a beam of different states in one call; per state the plan is the reference's, and the order of the
    plan is deterministic (two calls, same list)."""
    from billiard_ai_b200 import Engine, workloads as W
    from billiard_ai_b200.workloads import SNOOKER_SCORES
    eng = Engine.from_table(table, SNOOKER_SCORES, cue_type=TYPE_CODE["WHITE"])
    try:
        for vm in (1, 0):
            eng.set_planner(table, velocity_multiply=bool(vm))
            nb = 16
            layouts = [W.random_layout(nb, 200 + s) for s in range(24)]
            pos = np.stack([p for p, _ in layouts])
            types = np.stack([t for _, t in layouts])
            base = (np.arange(len(layouts), dtype=np.uint64) * np.uint64(777))
            eng.set_states(pos, types, 1 << TYPE_CODE["RED"], base_cost=base)
            n = eng.plan()
            plan = eng.fetch_plan()
            assert n == len(plan["vx"]) and n > 0 and (np.diff(plan["state"]) >= 0).all()
            again = (eng.plan(), eng.fetch_plan())
            assert again[0] == n and all(np.array_equal(plan[k], again[1][k]) for k in plan)
            for s in range(len(layouts)):
                want = _reference_plan(check, pos[s], types[s], np.array([TYPE_CODE["RED"]], np.int32), vm, base=int(base[s]))
                m = plan["state"] == s
                assert _multiset(plan["vx"][m], plan["vy"][m], plan["search_cost"][m], plan["cost"][m]) == \
                    _multiset(want["vx"], want["vy"], want["search_cost"], want["cost"]), (vm, s)
    finally:
        eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["test_01", "situation1", "situation2"])
@pytest.mark.parametrize("vm", [0, 1])
def test_device_plan_and_simulation_give_the_reference_solution_costs(table, ref, name, vm):
    """Config 1 without a host planner: bcu_plan + bcu_evaluate_plan on the reference's layouts must reproduce the reference's
    ranked solution costs (ref_plan: expand() best-first to exhaustion with the CPU simulation) -- incl. the thesis totals."""
    from billiard_ai_b200 import Engine
    from billiard_ai_b200.workloads import SNOOKER_SCORES
    pos, types, searched = _layout(name)
    mask = 0
    for t in searched:
        mask |= 1 << int(t)
    ref.set_toggles(0, vm)
    pl = ref.plan(pos, types, mask)
    want = sorted(int(c) for c in pl["result_cost"] if c != NONE)
    eng = Engine.from_table(table, SNOOKER_SCORES, cue_type=TYPE_CODE["WHITE"])
    try:
        eng.set_planner(table, velocity_multiply=bool(vm))
        eng.set_states(pos, types, mask)
        n = eng.plan()
        assert n == len(pl["vx"])
        out = eng.evaluate_plan(final_pos=False)
        got = sorted(int(c) for c, f in zip(out["cost"], out["flags"]) if f & 2)
        assert got == want and len(got) > 0
        tc, ti = eng.topk(min(64, len(got)))
        assert tc.tolist() == want[:len(tc)]
        if name == "situation1":
            with open(os.path.join(GOLDEN, "thesis_costs.json")) as f:
                thesis = json.load(f)["lists"]["s1_three_speeds" if vm else "s1_with_banks"]
            assert got[:len(thesis)] == thesis
    finally:
        eng.close()
    ref.set_toggles(0, 0)


@pytest.mark.gpu
def test_two_break_search_with_the_device_planner_gives_the_thesis_list(table):
    """bcu_search_levels in planner mode (n_dirs = 0): per break the beam is planned AND simulated on the device, the rules between
    the breaks run in the library's C++ pass -- no host planner, no Python in the loop.  With a beam wider than the valid strikes
    of a break the level-synchronous search finds what the reference's best-first search converges to: Situation 2 with the
    snooker rules gives the two-break totals printed in the thesis (183281, 185129, ...; tests/test_oracle_ref.py pins the same
    list on the reference itself), Situation 1 with one break the single-break list."""
    from billiard_ai_b200 import Engine
    from billiard_ai_b200.rules import SingleBreakRules, SnookerRules
    from billiard_ai_b200.search import search_levels
    from billiard_ai_b200.workloads import SNOOKER_SCORES
    with open(os.path.join(GOLDEN, "thesis_costs.json")) as f:
        thesis = json.load(f)["lists"]
    eng = Engine.from_table(table, SNOOKER_SCORES, cue_type=TYPE_CODE["WHITE"])
    try:
        eng.set_planner(table, velocity_multiply=False)
        pos, types, searched = _layout("situation2")
        info, stats = {}, []
        sols = search_levels(eng, pos, types, searched, 2, 1024, None, None, SnookerRules(), info=info, stats=stats)
        costs = [s["cost"] for s in sols]
        print(info, stats, costs[:8])
        assert costs[:len(thesis["s2_two_breaks"])] == thesis["s2_two_breaks"]
        assert all(s["breaks"] == 2 and len(s["strikes"]) == 2 and s["dirs"] == [-1, -1] for s in sols)
        # every break's strike re-simulates to the frame the next break was planned on (trace of the first solution)
        nl, times, rec = eng.simulate_trace(sols[0]["frames"][0][0], sols[0]["frames"][0][1], *sols[0]["strikes"][0])
        assert nl > 1
        pos1, types1, searched1 = _layout("situation1")
        one = search_levels(eng, pos1, types1, searched1, 1, 1024, None, None, SingleBreakRules())
        assert [s["cost"] for s in one][:len(thesis["s1_with_banks"])] == thesis["s1_with_banks"]
    finally:
        eng.close()
