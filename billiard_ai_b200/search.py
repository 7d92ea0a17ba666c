"""Multi-break search on top of the engine (BASELINE configs 3 and 4): a ctypes binding of bcu_search_levels
(include/billiard_cuda.h; csrc/bcu_search.cpp), the level-synchronous restatement of the reference's best-first search over
breaks (billiard_search.cpp:74-98, expand :224-244, next-break seeding :1407-1462):

    level L:  beam states x (directions x speeds)  --engine-->  cost = state's accumulated cost + simulationCost,
              validity against the state's expected types      --device top-k on (cost, id), merged across GPUs-->
              the k best strikes with their end states         --rules, one C++ pass over the k records-->
              level L + 1 states / solutions

The whole loop runs inside the library (C++); nothing here touches a candidate.  Every rank calls it with the same
arguments; the candidate ids of a level are sharded by contiguous range over the ranks of the engine's communicator and
the per-rank beams are merged by the engine's one collective, after which all ranks hold the identical beam and apply the
rules identically -- no other exchange.
"""
import ctypes as C

import numpy as np

from .rules import BreakRulesDesc
from .workloads import GRID_SEARCH_COST


class _SearchParams(C.Structure):
    _fields_ = [("levels", C.c_int32), ("beam", C.c_int32), ("n_dirs", C.c_int32), ("n_speeds", C.c_int32),
                ("dirs", C.POINTER(C.c_float)), ("speeds", C.POINTER(C.c_float)), ("search_cost", C.c_uint64),
                ("n_search_types", C.c_int32), ("search_types", C.POINTER(C.c_int32)), ("min_solutions", C.c_int32),
                ("max_duration_ms", C.c_uint32), ("cancel", C.POINTER(C.c_int32))]


class _LevelStats(C.Structure):
    _fields_ = [("states", C.c_int32), ("beam", C.c_int32), ("candidates", C.c_int64), ("kernel_ms", C.c_float),
                ("rules_ms", C.c_float), ("wall_ms", C.c_float)]


class _SearchResult(C.Structure):
    _fields_ = [("capacity", C.c_int32), ("count", C.c_int32), ("total", C.c_int32), ("levels_run", C.c_int32),
                ("stopped", C.c_int32), ("wall_ms", C.c_float), ("cost", C.POINTER(C.c_uint64)), ("breaks", C.POINTER(C.c_int32)),
                ("dir_index", C.POINTER(C.c_int32)), ("speed_index", C.POINTER(C.c_int32)), ("strike", C.POINTER(C.c_float)),
                ("frame_pos", C.POINTER(C.c_float)), ("frame_types", C.POINTER(C.c_int32)), ("end_pos", C.POINTER(C.c_float)),
                ("end_present", C.POINTER(C.c_uint8)), ("stats", C.POINTER(_LevelStats))]


STOPPED = {0: "exhausted", 1: "enough_solutions", 2: "timeout", 3: "cancelled"}


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def search_levels(engine, pos, types, search_types, levels, beam, dirs, speeds, rules, search_cost=GRID_SEARCH_COST,
                  rank=None, world=None, stats=None, min_solutions=0, max_duration_ms=0, cancel=None, info=None):
    """-> solutions, ascending by (cost, discovery order): dicts with `cost`, `strikes` (one (vx, vy) per break),
    `dirs` / `speeds` (grid indices per break), `frames` (the (pos, types) frame each break was played on), `pos` /
    `present` (table after the last break, rules applied when a further break was seeded) and `breaks`.

    pos, types: the frame (types < 0: no ball in that slot).  search_types: type codes the first break may pot.
    rules: SnookerRules / RepeatRules / SingleBreakRules (rules.py) -- handed to the library as a bcu_break_rules descriptor.
    `stats`, if given, receives per-level dicts (states, candidates, beam entries, ms of device / rules / wall time);
    `info` (a dict) receives `stopped`, `total`, `levels_run`, `wall_ms`.  rank / world are those of the engine's
    communicator (bcu_comm_init); the arguments are only checked against it.
    min_solutions / max_duration_ms / cancel (a ctypes c_int32 another thread may set): the three ways the reference's
    manager ends a search (process.hpp:188-198), looked at between levels."""
    pos = np.ascontiguousarray(pos, np.float32)
    types = np.ascontiguousarray(types, np.int32)
    planned = dirs is None   # the candidates of a state are the strikes the reference's planner finds (engine.set_planner)
    dirs = np.ascontiguousarray(dirs if not planned else np.zeros((0, 2)), np.float32)
    speeds = np.ascontiguousarray(speeds if not planned else np.zeros(0), np.float32)
    st = np.ascontiguousarray(search_types, np.int32)
    nb, cap = len(types), int(levels) * int(beam)
    if rank is not None or world is not None:
        er, ew = engine.rank_info()
        assert (rank is None or rank == er) and (world is None or world == ew), "rank / world differ from the engine's communicator"
    desc = rules if isinstance(rules, BreakRulesDesc) else rules.descriptor()
    sp = _SearchParams(int(levels), int(beam), len(dirs), len(speeds), _p(dirs, C.c_float), _p(speeds, C.c_float),
                       int(search_cost), len(st), _p(st, C.c_int32), int(min_solutions), int(max_duration_ms),
                       C.pointer(cancel) if cancel is not None else C.POINTER(C.c_int32)())
    cost = np.zeros(cap, np.uint64); breaks = np.zeros(cap, np.int32)
    di = np.zeros((cap, levels), np.int32); si = np.zeros((cap, levels), np.int32)
    strike = np.zeros((cap, levels, 2), np.float32)
    fpos = np.zeros((cap, levels, nb, 2), np.float32); ftypes = np.zeros((cap, levels, nb), np.int32)
    epos = np.zeros((cap, nb, 2), np.float32); epresent = np.zeros((cap, nb), np.uint8)
    lst = (_LevelStats * int(levels))()
    res = _SearchResult(cap, 0, 0, 0, 0, 0.0, _p(cost, C.c_uint64), _p(breaks, C.c_int32), _p(di, C.c_int32), _p(si, C.c_int32),
                        _p(strike, C.c_float), _p(fpos, C.c_float), _p(ftypes, C.c_int32), _p(epos, C.c_float),
                        _p(epresent, C.c_uint8), lst)
    engine._ck(engine.lib.bcu_search_levels(engine.h, C.c_int32(nb), _p(pos, C.c_float), _p(types, C.c_int32), C.byref(sp),
                                            C.byref(desc), C.byref(res)), "bcu_search_levels")
    if stats is not None:
        for lv in range(res.levels_run):
            s = lst[lv]
            stats.append(dict(level=lv, states=s.states, candidates=s.candidates, beam=s.beam, kernel_ms=s.kernel_ms,
                              rules_ms=s.rules_ms, wall_ms=s.wall_ms))
    if info is not None:
        info.update(stopped=STOPPED.get(res.stopped, res.stopped), total=res.total, levels_run=res.levels_run, wall_ms=res.wall_ms)
    out = []
    for r in range(res.count):
        b = int(breaks[r])
        out.append(dict(cost=int(cost[r]), breaks=b, strikes=[(float(strike[r, l, 0]), float(strike[r, l, 1])) for l in range(b)],
                        dirs=di[r, :b].tolist(), speeds=si[r, :b].tolist(), frames=[(fpos[r, l], ftypes[r, l]) for l in range(b)],
                        pos=epos[r], present=epresent[r]))
    return out


def solution_traces(engine, solutions, top=None):
    """Full per-layer traces of every break of the first `top` solutions -- what billiard::search::search() returns as one
    node::System per break of each solution (billiard_search.hpp:9-12) -- in ONE device call
    (bcu_trace_strikes).  -> per solution a list, per break a dict `layers`, `times [layers]`, `rec [layers][balls][10]`
    (record layout: include/billiard_cuda.h, bcu_simulate_trace)."""
    sols = solutions if top is None else solutions[:top]
    if not sols:
        return []
    frames, slot, sidx, vx, vy = [], {}, [], [], []
    for s in sols:
        for (fpos, ftypes), (svx, svy) in zip(s["frames"], s["strikes"]):
            key = (fpos.tobytes(), ftypes.tobytes())
            if key not in slot:
                slot[key] = len(frames)
                frames.append((fpos, ftypes))
            sidx.append(slot[key]); vx.append(svx); vy.append(svy)
    engine.set_states(np.stack([f[0] for f in frames]), np.stack([f[1] for f in frames]), 0xFFFFFFFF)
    layers, times, rec = engine.trace_strikes(np.array(sidx, np.int32), np.array(vx, np.float32), np.array(vy, np.float32))
    out, i = [], 0
    for s in sols:
        per = []
        for _ in s["strikes"]:
            n = int(layers[i])
            per.append(dict(layers=n, times=times[i, :n].copy(), rec=rec[i, :n].copy()))
            i += 1
        out.append(per)
    return out
