"""Rules between breaks: ctypes binding of bcu_snooker_next_search / bcu_snooker_state_after_break
(include/billiard_cuda.h; reference: billiard::snooker::nextSearch / stateAfterBreak,
lib/billiard_snooker/src/billiard_snooker.cpp:792-822, 1097-1126), plus the data-driven rule sets the level driver
(search.py) takes.  Host logic only -- the reference runs these on the CPU once per finished simulation as well."""
import ctypes as C

import numpy as np

from .engine import BcuError, load_library
from .workloads import SNOOKER_SPOTS, TYPE_CODE

COLOR_NAMES = ("YELLOW", "GREEN", "BROWN", "BLUE", "PINK", "BLACK")   # typesToReplace, billiard_snooker.cpp:773-780


class _SnookerConfig(C.Structure):
    _fields_ = [("red_type", C.c_int32), ("color_types", C.c_int32 * 6), ("spots", (C.c_float * 2) * 6),
                ("head_rail_dir", C.c_float * 2), ("diameter_squared", C.c_float), ("radius", C.c_float),
                ("minimum", C.c_float * 2), ("maximum", C.c_float * 2)]


class BreakRulesDesc(C.Structure):
    """bcu_break_rules (include/billiard_cuda.h): the rules between breaks as data for bcu_search_levels."""
    _fields_ = [("kind", C.c_int32), ("snooker", _SnookerConfig)]


BREAK_RULES_SINGLE, BREAK_RULES_REPEAT, BREAK_RULES_SNOOKER = 0, 1, 2


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class SnookerRules:
    """billiard::snooker rules for the multi-break search.

    Defaults are the reference's own configuration: spots of billiard_ui/Assets/Resources/configuration.json, head rail
    direction (1, 0), radius 26.15 mm, table 1881 x 943 mm (test/billiard_snooker/src/billiard_snooker_test.cpp:288-299).
    """

    name = "snooker"

    def __init__(self, spots=None, head_rail_dir=(1.0, 0.0), radius=26.15, minimum=(-1881.0 / 2.0, -943.0 / 2.0),
                 maximum=(1881.0 / 2.0, 943.0 / 2.0)):
        self.lib = load_library()
        spots = spots or SNOOKER_SPOTS
        c = _SnookerConfig()
        c.red_type = TYPE_CODE["RED"]
        for k, nme in enumerate(COLOR_NAMES):
            c.color_types[k] = TYPE_CODE[nme]
            c.spots[k][0] = np.float32(spots[nme][0]); c.spots[k][1] = np.float32(spots[nme][1])
        c.head_rail_dir[0], c.head_rail_dir[1] = np.float32(head_rail_dir[0]), np.float32(head_rail_dir[1])
        r = np.float32(radius)
        c.radius = r
        c.diameter_squared = np.float32(np.float32(r * np.float32(2.0)) * np.float32(r * np.float32(2.0)))
        c.minimum[0], c.minimum[1] = np.float32(minimum[0]), np.float32(minimum[1])
        c.maximum[0], c.maximum[1] = np.float32(maximum[0]), np.float32(maximum[1])
        self.cfg = c
        self.spots = np.array([[c.spots[k][0], c.spots[k][1]] for k in range(6)], np.float32)

    def descriptor(self):
        d = BreakRulesDesc()
        d.kind = BREAK_RULES_SNOOKER
        d.snooker = self.cfg
        return d

    def next_search(self, types, present, previous_types):
        """-> list of type codes the next break searches for (-1 = the reference's empty type name)."""
        types = np.ascontiguousarray(types, np.int32)
        present = np.ascontiguousarray(present, np.uint8)
        prev = np.ascontiguousarray(previous_types, np.int32)
        out = np.zeros(8, np.int32)
        cnt = C.c_int32(0)
        rc = self.lib.bcu_snooker_next_search(C.byref(self.cfg), C.c_int32(len(types)), _p(types, C.c_int32),
                                              _p(present, C.c_uint8), C.c_int32(len(prev)), _p(prev, C.c_int32),
                                              _p(out, C.c_int32), C.byref(cnt))
        if rc != 0:
            raise BcuError(f"bcu_snooker_next_search: {rc}")
        return out[:cnt.value].tolist()

    def state_after_break(self, pos, types, present):
        """-> (ok, pos', present'): potted colours back on the table while a red is left; ok False = the reference's
        empty State (a colour could not be placed: the candidate is dropped)."""
        pos = np.ascontiguousarray(pos, np.float32)
        types = np.ascontiguousarray(types, np.int32)
        present = np.ascontiguousarray(present, np.uint8)
        out_pos = np.zeros_like(pos)
        out_present = np.zeros_like(present)
        ok = C.c_int32(0)
        rc = self.lib.bcu_snooker_state_after_break(C.byref(self.cfg), C.c_int32(len(types)), _p(pos, C.c_float),
                                                    _p(types, C.c_int32), _p(present, C.c_uint8), _p(out_pos, C.c_float),
                                                    _p(out_present, C.c_uint8), C.byref(ok))
        if rc != 0:
            raise BcuError(f"bcu_snooker_state_after_break: {rc}")
        return bool(ok.value), out_pos, out_present


class RepeatRules:
    """Not in the reference (it ships snooker rules only; Configuration's default rules end the search after one break,
    type.hpp:324-332): the next break searches for the same types again and the state is left as it is.  Used for the
    pool layouts of BASELINE configs (rules as a plug-in).  The search itself runs the C++ restatement
    (BCU_BREAK_RULES_REPEAT in csrc/bcu_search.cpp); the methods below are the host-side statement of the same rule."""

    name = "repeat"

    def descriptor(self):
        d = BreakRulesDesc()
        d.kind = BREAK_RULES_REPEAT
        return d

    def next_search(self, types, present, previous_types):
        return list(previous_types)

    def state_after_break(self, pos, types, present):
        return True, np.array(pos, np.float32, copy=True), np.array(present, np.uint8, copy=True)


class SingleBreakRules:
    """The reference Configuration's default rules (type.hpp:325-328): nextSearch returns an empty Search, so every valid
    strike is a solution after one break; the state is left as it is."""

    name = "single"

    def descriptor(self):
        d = BreakRulesDesc()
        d.kind = BREAK_RULES_SINGLE
        return d

    def next_search(self, types, present, previous_types):
        return []

    def state_after_break(self, pos, types, present):
        return True, np.array(pos, np.float32, copy=True), np.array(present, np.uint8, copy=True)
