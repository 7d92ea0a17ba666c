"""ctypes bindings for the two oracles (TEST INFRASTRUCTURE ONLY).

  * RefOracle  -> oracle/_ref/libref_oracle.so : the reference's own sources compiled from
                  /root/reference by oracle/Makefile (prebuilt .so travels to the GPU box)
  * PortOracle -> oracle/_ref/libport_oracle.so: the plain C++ restatement in oracle/port/

Nothing under billiard_ai_b200/ imports this module.
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libref_oracle.so")
PORT_SO = os.path.join(ORACLE_DIR, "_ref", "libport_oracle.so")
GOLDEN = os.path.join(ROOT, "tests", "golden")

F_COMPLETED, F_VALID, F_OVERFLOW, F_NOEVENT = 1, 2, 4, 8
MAX_LAYERS = 32
REC = 10

# type codes shared by every test / bench in this repo
TYPE_NAMES = ["WHITE", "RED", "YELLOW", "GREEN", "BROWN", "BLUE", "PINK", "BLACK"]
TYPE_CODE = {n: i for i, n in enumerate(TYPE_NAMES)}
# billiard_snooker.cpp:751-759, 787-790: points / 7.0
SNOOKER_SCORES = np.array([0.0, 1 / 7.0, 2 / 7.0, 3 / 7.0, 4 / 7.0, 5 / 7.0, 6 / 7.0, 7 / 7.0], dtype=np.float64)


def build_oracles():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "all"])


def _ptr(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def load_table(path=None):
    with open(path or os.path.join(GOLDEN, "table.json")) as f:
        return json.load(f)


def table_arrays(tbl):
    rails6 = np.array([[r["start"][0], r["start"][1], r["end"][0], r["end"][1], r["shift"][0], r["shift"][1]]
                       for r in tbl["rails"]], dtype=np.float32)
    pockets7 = np.array([[p["position"][0], p["position"][1], p["radius"], p["normal"][0], p["normal"][1],
                          p["pottingPoint"][0], p["pottingPoint"][1]] for p in tbl["pockets"]], dtype=np.float32)
    return rails6, pockets7


class RefOracle:
    def __init__(self, tbl=None, scores=SNOOKER_SCORES):
        if not os.path.exists(REF_SO):
            build_oracles()
        self.lib = C.CDLL(REF_SO)
        self.tbl = tbl or load_table()
        rails6, pockets7 = table_arrays(self.tbl)
        self.R, self.P = len(rails6), len(pockets7)
        rail_ids = "\n".join(r["id"] for r in self.tbl["rails"]).encode()
        pocket_ids = "\n".join(p["id"] for p in self.tbl["pockets"]).encode()
        rc = self.lib.ref_set_table(C.c_int(self.R), rail_ids, _ptr(rails6, C.c_float), C.c_int(self.P), pocket_ids,
                                    _ptr(pockets7, C.c_float), C.c_float(np.float32(self.tbl["ballRadius"])),
                                    C.c_float(self.tbl["minimalPocketVelocity"]),
                                    C.c_float(self.tbl["innerTableLength"]), C.c_float(self.tbl["innerTableWidth"]))
        assert rc == 0
        sc = np.ascontiguousarray(scores, dtype=np.float64)
        self.lib.ref_set_rules("\n".join(TYPE_NAMES).encode(), _ptr(sc, C.c_double))
        self.lib.ref_plan.restype = C.c_int64

    def rail_derived(self):
        out = np.zeros((self.R, 6), dtype=np.float32)
        self.lib.ref_get_rail_derived(_ptr(out, C.c_float))
        return out  # normal.xy, shiftedStart.xy, shiftedEnd.xy

    def set_toggles(self, depth, velocity_multiply):
        self.lib.ref_set_toggles(C.c_int(int(depth)), C.c_int(int(velocity_multiply)))

    def simulate_trace(self, pos, types, vx, vy):
        pos = np.ascontiguousarray(pos, dtype=np.float32)
        types = np.ascontiguousarray(types, dtype=np.int32)
        n = len(types)
        times = np.zeros(MAX_LAYERS, dtype=np.float32)
        rec = np.zeros((MAX_LAYERS, n, REC), dtype=np.float32)
        nl = self.lib.ref_simulate_trace(C.c_int(n), _ptr(pos, C.c_float), _ptr(types, C.c_int32),
                                         C.c_float(vx), C.c_float(vy), C.c_int(MAX_LAYERS), _ptr(times, C.c_float),
                                         _ptr(rec, C.c_float))
        return nl, times[:nl], rec[:nl]

    def evaluate(self, state_pos, state_types, expected_mask, cand_state, vx, vy, search_cost, base_cost, threads=1,
                 event_log=False):
        return _evaluate(self.lib.ref_evaluate, state_pos, state_types, expected_mask, cand_state, vx, vy, search_cost,
                         base_cost, threads, event_log, with_gap=False)

    def plan(self, pos, types, expected_mask, max_out=4096):
        pos = np.ascontiguousarray(pos, dtype=np.float32)
        types = np.ascontiguousarray(types, dtype=np.int32)
        vx = np.zeros(max_out, np.float32); vy = np.zeros(max_out, np.float32)
        sc = np.zeros(max_out, np.uint64); bc = np.zeros(max_out, np.uint64); rc = np.zeros(max_out, np.uint64)
        sid = np.zeros(max_out, np.int32)
        exp = C.c_int64(0)
        n = self.lib.ref_plan(C.c_int(len(types)), _ptr(pos, C.c_float), _ptr(types, C.c_int32),
                              C.c_uint32(expected_mask), C.c_int64(max_out), _ptr(vx, C.c_float), _ptr(vy, C.c_float),
                              _ptr(sc, C.c_uint64), _ptr(bc, C.c_uint64), _ptr(rc, C.c_uint64), _ptr(sid, C.c_int32),
                              C.byref(exp))
        n = min(n, max_out)
        return dict(vx=vx[:n], vy=vy[:n], search_cost=sc[:n], base_cost=bc[:n], result_cost=rc[:n], sim_id=sid[:n],
                    expansions=exp.value)

    def quartic(self, a, b, c, d, e):
        out = np.zeros(8, dtype=np.float64)
        n = self.lib.ref_quartic(C.c_double(a), C.c_double(b), C.c_double(c), C.c_double(d), C.c_double(e),
                                 _ptr(out, C.c_double))
        return [complex(out[2 * i], out[2 * i + 1]) for i in range(n)]

    def quartic_batch(self, coef):
        coef = np.ascontiguousarray(coef, dtype=np.float64)
        n = len(coef)
        out = np.zeros((n, 4, 2), dtype=np.float64)
        cnt = np.zeros(n, dtype=np.int32)
        self.lib.ref_quartic_batch(C.c_int64(n), _ptr(coef, C.c_double), _ptr(out, C.c_double), _ptr(cnt, C.c_int32))
        return out, cnt

    def time_to_collision(self, a1, v1, p1, a2, v2, p2, diameter):
        arr = [np.array(x, dtype=np.float32) for x in (a1, v1, p1, a2, v2, p2)]
        t = C.c_float(0)
        has = self.lib.ref_time_to_collision(*[_ptr(x, C.c_float) for x in arr], C.c_float(diameter), C.byref(t))
        return (t.value if has else None)

    def time_to_collision_batch(self, p1, v1, r1, p2, v2, r2, diameter):
        return _ttc_batch(self.lib.ref_time_to_collision_batch, p1, v1, r1, p2, v2, r2, diameter)

    def initial_velocity(self, dx, dy, frames):
        out = np.zeros(2, dtype=np.float32)
        self.lib.ref_calc_initial_velocity(C.c_float(dx), C.c_float(dy), C.c_int(frames), _ptr(out, C.c_float))
        return float(out[0]), float(out[1])


def _ttc_batch(fn, p1, v1, r1, p2, v2, r2, diameter):
    p1, v1, p2, v2 = [np.ascontiguousarray(x, dtype=np.float32) for x in (p1, v1, p2, v2)]
    r1, r2 = [np.ascontiguousarray(x, dtype=np.int32) for x in (r1, r2)]
    n = len(r1)
    t = np.zeros(n, dtype=np.float32)
    has = np.zeros(n, dtype=np.int32)
    fn(C.c_int64(n), _ptr(p1, C.c_float), _ptr(v1, C.c_float), _ptr(r1, C.c_int32), _ptr(p2, C.c_float),
       _ptr(v2, C.c_float), _ptr(r2, C.c_int32), C.c_float(diameter), _ptr(t, C.c_float), _ptr(has, C.c_int32))
    return t, has


def _evaluate(fn, state_pos, state_types, expected_mask, cand_state, vx, vy, search_cost, base_cost, threads,
              event_log, with_gap):
    state_pos = np.ascontiguousarray(state_pos, dtype=np.float32)
    state_types = np.ascontiguousarray(state_types, dtype=np.int32)
    ns, n = state_types.shape
    expected_mask = np.ascontiguousarray(np.broadcast_to(np.asarray(expected_mask, dtype=np.uint32), (ns,)))
    vx = np.ascontiguousarray(vx, dtype=np.float32)
    vy = np.ascontiguousarray(vy, dtype=np.float32)
    nc = len(vx)
    cand_state = np.ascontiguousarray(np.broadcast_to(np.asarray(cand_state, dtype=np.int32), (nc,)))
    search_cost = np.ascontiguousarray(np.broadcast_to(np.asarray(search_cost, dtype=np.uint64), (nc,)))
    base_cost = np.ascontiguousarray(np.broadcast_to(np.asarray(base_cost, dtype=np.uint64), (nc,)))
    cost = np.zeros(nc, np.uint64); flags = np.zeros(nc, np.uint32); nev = np.zeros(nc, np.int32)
    potted = np.zeros(nc, np.uint32); ftime = np.zeros(nc, np.float32); fpos = np.zeros((nc, n, 2), np.float32)
    log = np.full((nc, 32, 4), -1, np.int32) if event_log else None
    args = [C.c_int(ns), C.c_int(n), _ptr(state_pos, C.c_float), _ptr(state_types, C.c_int32),
            _ptr(expected_mask, C.c_uint32), C.c_int64(nc), _ptr(cand_state, C.c_int32), _ptr(vx, C.c_float),
            _ptr(vy, C.c_float), _ptr(search_cost, C.c_uint64), _ptr(base_cost, C.c_uint64), C.c_int(threads),
            _ptr(cost, C.c_uint64), _ptr(flags, C.c_uint32), _ptr(nev, C.c_int32), _ptr(potted, C.c_uint32),
            _ptr(ftime, C.c_float), _ptr(fpos, C.c_float),
            _ptr(log, C.c_int32) if event_log else C.POINTER(C.c_int32)()]
    gap = None
    if with_gap:
        gap = np.zeros(nc, np.float32)
        args.append(_ptr(gap, C.c_float))
    fn(*args)
    return dict(cost=cost, flags=flags, n_events=nev, potted=potted, final_time=ftime, final_pos=fpos, event_log=log,
                min_gap=gap)


class PortOracle:
    def __init__(self, tbl=None, scores=SNOOKER_SCORES, rail_derived=None):
        if not os.path.exists(PORT_SO):
            build_oracles()
        self.lib = C.CDLL(PORT_SO)
        self.tbl = tbl or load_table()
        if rail_derived is None:
            rail_derived = np.array(self.tbl["railsDerived"], dtype=np.float32)  # normal, shiftedStart, shiftedEnd
        rd = np.ascontiguousarray(rail_derived, dtype=np.float32)
        # port order: shiftedStart, shiftedEnd, normal
        r6 = np.ascontiguousarray(np.concatenate([rd[:, 2:4], rd[:, 4:6], rd[:, 0:2]], axis=1), dtype=np.float32)
        _, pockets7 = table_arrays(self.tbl)
        p3 = np.ascontiguousarray(pockets7[:, :3])
        diameter = np.float32(np.float32(self.tbl["ballRadius"]) * np.float32(2))
        self.lib.port_set_table(C.c_int(len(r6)), _ptr(r6, C.c_float), C.c_int(len(p3)), _ptr(p3, C.c_float),
                                C.c_float(diameter))
        sc = np.ascontiguousarray(scores, dtype=np.float64)
        self.lib.port_set_rules(C.c_int(len(sc)), C.c_int(TYPE_CODE["WHITE"]), _ptr(sc, C.c_double))

    def simulate_trace(self, pos, types, vx, vy):
        pos = np.ascontiguousarray(pos, dtype=np.float32)
        types = np.ascontiguousarray(types, dtype=np.int32)
        n = len(types)
        times = np.zeros(MAX_LAYERS, dtype=np.float32)
        rec = np.zeros((MAX_LAYERS, n, REC), dtype=np.float32)
        gap = C.c_float(0)
        nl = self.lib.port_simulate_trace(C.c_int(n), _ptr(pos, C.c_float), _ptr(types, C.c_int32), C.c_float(vx),
                                          C.c_float(vy), C.c_int(MAX_LAYERS), _ptr(times, C.c_float),
                                          _ptr(rec, C.c_float), C.byref(gap))
        return nl, times[:nl], rec[:nl], gap.value

    def evaluate(self, state_pos, state_types, expected_mask, cand_state, vx, vy, search_cost, base_cost, threads=1,
                 event_log=False):
        return _evaluate(self.lib.port_evaluate, state_pos, state_types, expected_mask, cand_state, vx, vy, search_cost,
                         base_cost, threads, event_log, with_gap=True)

    def quartic(self, a, b, c, d, e):
        out = np.zeros(8, dtype=np.float64)
        n = self.lib.port_quartic(C.c_double(a), C.c_double(b), C.c_double(c), C.c_double(d), C.c_double(e),
                                  _ptr(out, C.c_double))
        return [complex(out[2 * i], out[2 * i + 1]) for i in range(n)]

    def time_to_collision_batch(self, p1, v1, r1, p2, v2, r2, diameter):
        return _ttc_batch(self.lib.port_time_to_collision_batch, p1, v1, r1, p2, v2, r2, diameter)

    def counters(self, reset=True):
        out = np.zeros(6, dtype=np.uint64)
        self.lib.port_get_counters(_ptr(out, C.c_uint64), C.c_int(int(reset)))
        return dict(zip(["events", "pair_tests", "quartics", "rail_tests", "pocket_tests", "near_ties"], out.tolist()))
