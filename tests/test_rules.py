"""Rules between breaks: the product's bcu_snooker_next_search / bcu_snooker_state_after_break
against (a) the assertions of the reference's own test file, test/billiard_snooker/src/billiard_snooker_test.cpp:20-281,
and (b) the reference's unmodified functions (oracle/_ref/libref_rules.so, built by oracle/Makefile from the lines of
billiard_snooker.cpp where they lie) on random frames.  Host logic only: runs without a GPU."""
import ctypes as C
import os

import numpy as np
import pytest

from billiard_ai_b200.rules import COLOR_NAMES, SnookerRules
from billiard_ai_b200.workloads import SNOOKER_SPOTS, TYPE_CODE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_RULES_SO = os.path.join(ROOT, "oracle", "_ref", "libref_rules.so")
T = TYPE_CODE
SPOT = {k: np.array(v, np.float32) for k, v in SNOOKER_SPOTS.items()}


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class RefRules:
    """The reference's own nextSearch / stateAfterBreak (test infrastructure)."""

    def __init__(self, rules):
        if not os.path.exists(REF_RULES_SO):
            import subprocess
            subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "rules"])
        self.lib = C.CDLL(REF_RULES_SO)
        c = rules.cfg
        spots = np.ascontiguousarray(rules.spots.reshape(-1), np.float32)
        head = np.array([c.head_rail_dir[0], c.head_rail_dir[1]], np.float32)
        mn = np.array([c.minimum[0], c.minimum[1]], np.float32); mx = np.array([c.maximum[0], c.maximum[1]], np.float32)
        self.lib.refrules_configure(_p(spots, C.c_float), _p(head, C.c_float), C.c_float(c.diameter_squared),
                                    C.c_float(c.radius), _p(mn, C.c_float), _p(mx, C.c_float))

    def next_search(self, pos, types, present, previous):
        pos = np.ascontiguousarray(pos, np.float32); types = np.ascontiguousarray(types, np.int32)
        present = np.ascontiguousarray(present, np.uint8); prev = np.ascontiguousarray(previous, np.int32)
        out = np.zeros(8, np.int32)
        k = self.lib.refrules_next_search(C.c_int(len(types)), _p(pos, C.c_float), _p(types, C.c_int32), _p(present, C.c_uint8),
                                          C.c_int(len(prev)), _p(prev, C.c_int32), _p(out, C.c_int32))
        return out[:k].tolist()

    def state_after_break(self, pos, types, present):
        pos = np.ascontiguousarray(pos, np.float32); types = np.ascontiguousarray(types, np.int32)
        present = np.ascontiguousarray(present, np.uint8)
        op = np.zeros_like(pos); opr = np.zeros_like(present)
        ok = self.lib.refrules_state_after_break(C.c_int(len(types)), _p(pos, C.c_float), _p(types, C.c_int32),
                                                 _p(present, C.c_uint8), _p(op, C.c_float), _p(opr, C.c_uint8))
        return bool(ok), op, opr


@pytest.fixture(scope="module")
def rules():
    return SnookerRules()


@pytest.fixture(scope="module")
def ref(rules):
    return RefRules(rules)


def _frame(balls):
    """balls: list of (type name, (x, y) or None for a potted ball) -> pos, types, present"""
    pos = np.array([b[1] if b[1] is not None else (0.0, 0.0) for b in balls], np.float32).reshape(-1, 2)
    types = np.array([T[b[0]] for b in balls], np.int32)
    present = np.array([b[1] is not None for b in balls], np.uint8)
    return pos, types, present


# ---- nextSearch: billiard_snooker_test.cpp:20-97 ---------------------------------------------------------------------
def test_next_search_reference_cases(rules, ref):
    all_colors = [T[c] for c in ("BLACK", "PINK", "BLUE", "BROWN", "GREEN", "YELLOW")]
    one_red = _frame([("RED", (0, 0))])
    empty = (np.zeros((0, 2), np.float32), np.zeros(0, np.int32), np.zeros(0, np.uint8))
    cases = [(one_red, [T["RED"]], sorted(all_colors)),                                   # :20-37
             (empty, [T["RED"]], [T["YELLOW"]]),                                          # :39-48
             (one_red, [T[c] for c in ("BLACK", "PINK", "BROWN", "BLUE", "GREEN", "YELLOW")], [T["RED"]])]   # :50-62
    for a, b in (("YELLOW", "GREEN"), ("GREEN", "BROWN"), ("BROWN", "BLUE"), ("BLUE", "PINK"), ("PINK", "BLACK")):   # :88-97
        cases.append((empty, [T[a]], [T[b]]))
    for (pos, types, present), prev, want in cases:
        got = rules.next_search(types, present, prev)
        assert sorted(got) == sorted(want)
        assert got == ref.next_search(pos, types, present, prev)
    assert rules.next_search(one_red[1], one_red[2], []) == []                            # billiard_snooker.cpp:796-798
    assert rules.next_search(empty[1], empty[2], [T["BLACK"]]) == [-1]                    # successor of BLACK: ""
    assert ref.next_search(*empty, [T["BLACK"]]) == [-1]


# ---- stateAfterBreak: billiard_snooker_test.cpp:126-281 --------------------------------------------------------------
def _check_state(rules, ref, frame):
    pos, types, present = frame
    ok, p2, pr2 = rules.state_after_break(pos, types, present)
    rok, rp2, rpr2 = ref.state_after_break(pos, types, present)
    assert ok == rok
    assert np.array_equal(pr2, rpr2)
    on = pr2.astype(bool)
    assert np.array_equal(p2[on].view(np.uint32), rp2[on].view(np.uint32))   # bit-identical positions
    return ok, p2, pr2


def test_state_after_break_reference_cases(rules, ref):
    # :126-141 nothing to do when every colour is on the table (and there is no red: state returned as is)
    frame = _frame([(c, (i, i)) for i, c in enumerate(COLOR_NAMES)])
    ok, p2, pr2 = _check_state(rules, ref, frame)
    assert ok and np.array_equal(p2, frame[0]) and pr2.all()
    # :146-180 a missing colour returns to its own spot while a red is on the table
    for missing in COLOR_NAMES:
        frame = _frame([("RED", (-1000, -1000))] + [(c, None if c == missing else tuple(SPOT[c])) for c in COLOR_NAMES])
        ok, p2, pr2 = _check_state(rules, ref, frame)
        assert ok and pr2.sum() == 7
        assert np.array_equal(p2[1 + COLOR_NAMES.index(missing)], SPOT[missing])
    # :182-235 own spot occupied by a red: the next higher FREE spot
    order = list(COLOR_NAMES)
    for i, missing in enumerate(order):
        for expected in order[i + 1:]:
            balls = []
            for c in COLOR_NAMES:
                if c == missing:
                    balls.append(("RED", tuple(SPOT[c])))
                elif c == expected:
                    balls.append((c, (-1000, -1000)))
                else:
                    balls.append((c, tuple(SPOT[c])))
            balls.append((missing, None))
            ok, p2, pr2 = _check_state(rules, ref, _frame(balls))
            assert ok and pr2.sum() == 7
            assert np.array_equal(p2[-1], SPOT[expected]), (missing, expected)
    # :237-257 / :260-281 every spot occupied: next to the own spot, within 5 radii, on the table
    max_d2 = np.float32(26.15 * 5) ** 2
    frame = _frame([("RED", tuple(SPOT["YELLOW"]))] + [(c, tuple(SPOT[c])) for c in COLOR_NAMES[1:]] + [("YELLOW", None)])
    ok, p2, pr2 = _check_state(rules, ref, frame)
    assert ok and pr2.sum() == 7 and ((p2[-1] - SPOT["YELLOW"]) ** 2).sum() <= max_d2
    frame = _frame([("YELLOW", tuple(SPOT["YELLOW"]))] + [(c, tuple(SPOT[c])) for c in COLOR_NAMES[1:5]] +
                   [("RED", tuple(SPOT["BLACK"])), ("BLACK", None)])
    ok, p2, pr2 = _check_state(rules, ref, frame)
    assert ok and pr2.sum() == 7 and ((p2[-1] - SPOT["BLACK"]) ** 2).sum() <= max_d2
    assert abs(p2[-1][0]) <= 1881.0 / 2 - 26.15 and abs(p2[-1][1]) <= 943.0 / 2 - 26.15


def test_state_after_break_random_frames_match_the_reference(rules, ref):
    """2000 random frames: reds crowding the spots (so that the chain of higher spots and the area search both run),
    random subsets of colours potted, sometimes no red at all; positions must be bit-identical to the reference's."""
    rng = np.random.RandomState(11)
    n_area = n_fail = 0
    spots = np.array([SPOT[c] for c in COLOR_NAMES])
    for it in range(2000):
        n_red = rng.randint(0, 16)
        balls = [("WHITE", tuple(rng.uniform(-800, 800, 2) * (1, 0.5)))]
        for _ in range(n_red):
            if rng.rand() < 0.6:   # on or near a spot
                s = spots[rng.randint(6)] + (rng.uniform(-60, 60, 2) if rng.rand() < 0.7 else 0.0)
                balls.append(("RED", tuple(s)))
            else:
                balls.append(("RED", tuple(rng.uniform(-900, 900, 2) * (1, 0.5))))
        for c in COLOR_NAMES:
            r = rng.rand()
            balls.append((c, None if r < 0.5 else tuple(SPOT[c]) if r < 0.75 else tuple(rng.uniform(-900, 900, 2) * (1, 0.5))))
        frame = _frame(balls)
        ok, p2, pr2 = _check_state(rules, ref, frame)
        n_fail += not ok
        if ok:
            newly = pr2.astype(bool) & ~frame[2].astype(bool)
            n_area += int(any(not (spots == p2[i]).all(axis=1).any() for i in np.nonzero(newly)[0]))
        prev = [T["RED"]] if rng.rand() < 0.5 else [T[c] for c in ("BLACK", "PINK", "BLUE", "BROWN", "GREEN", "YELLOW")]
        assert rules.next_search(frame[1], pr2, prev) == ref.next_search(p2, frame[1], pr2, prev)
    assert n_area > 20   # the area search was exercised
