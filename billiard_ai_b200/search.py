"""Level-synchronous multi-break search on top of the engine (BASELINE configs 3 and 4).

The reference searches break after break best-first (billiard_search.cpp:74-98, expand :1358-1471): a finished, valid
simulation either is a solution or -- when depth search is on and fewer than BREAKS simulations lie on its path -- seeds
the next break: the balls at rest form the next State, `_rules._nextSearch` names the types to pot next and
`_rules._modifyState` puts potted colours back (:1407-1462).  Here the same step is taken for a whole beam at once:

    level L:  beam states x (directions x speeds)  --engine-->  cost = state's accumulated cost + simulationCost,
              validity against the state's expected types      --device top-k on (cost, id), merged across GPUs-->
              the k best strikes with their end states         --rules (host, rules.py)-->  level L + 1 states / solutions

Every rank runs this loop; the candidate ids of a level are sharded by contiguous range and the per-rank beams are merged
by the engine's one collective (bcu_level_select), after which all ranks hold the identical beam and apply the rules
identically -- no other exchange.
"""
import numpy as np

from .workloads import GRID_SEARCH_COST, shard_range


def _mask_of(type_codes):
    m = 0
    for t in type_codes:
        if t >= 0:
            m |= 1 << int(t)
    return m


def search_levels(engine, pos, types, search_types, levels, beam, dirs, speeds, rules, search_cost=GRID_SEARCH_COST,
                  rank=0, world=1, stats=None):
    """-> solutions, ascending by (cost, discovery order): dicts with `cost`, `strikes` (one (vx, vy) per break),
    `dirs` / `speeds` (grid indices per break), `frames` (the (pos, types) frame each break was played on), `pos` /
    `present` (table after the last break, rules applied when a further break was seeded) and `breaks`.

    pos, types: the frame (types < 0: no ball in that slot).  search_types: type codes the first break may pot.
    `stats`, if given, receives per-level dicts (states, candidates, beam entries, ms of device time)."""
    pos = np.ascontiguousarray(pos, np.float32)
    types = np.ascontiguousarray(types, np.int32)
    n_dirs, n_speeds = len(dirs), len(speeds)
    per_state = n_dirs * n_speeds
    dirs = np.ascontiguousarray(dirs, np.float32)
    speeds = np.ascontiguousarray(speeds, np.float32)
    states = [dict(pos=pos, types=types, searched=[int(t) for t in search_types], cost=0, path=[])]
    solutions = []
    for level in range(levels):
        if not states:
            break
        s_pos = np.stack([s["pos"] for s in states])
        s_types = np.stack([s["types"] for s in states])
        s_mask = np.array([_mask_of(s["searched"]) for s in states], np.uint32)
        s_cost = np.array([s["cost"] for s in states], np.uint64)
        engine.set_states(s_pos, s_types, s_mask, base_cost=s_cost)
        total = len(states) * per_state
        lo, hi = shard_range(total, rank, world)
        engine.evaluate_grid(dirs, speeds, search_cost, first=lo, count=hi - lo)
        ms = engine.last_kernel_ms() if stats is not None else 0.0
        engine.level_select(beam)
        b = engine.fetch_beam()
        if stats is not None:
            stats.append(dict(level=level, states=len(states), candidates=total, beam=len(b["cost"]), kernel_ms=ms))
        nxt = []
        for cost, cid, potted, fpos in zip(b["cost"].tolist(), b["id"].tolist(), b["potted"].tolist(), b["final_pos"]):
            si, rem = divmod(int(cid), per_state)
            k, j = divmod(rem, n_speeds)
            parent = states[si]
            vx = np.float32(speeds[j] * dirs[k, 0]); vy = np.float32(speeds[j] * dirs[k, 1])
            path = parent["path"] + [(k, j, float(vx), float(vy), parent["pos"], parent["types"])]
            frame_types = parent["types"]
            present = ((frame_types >= 0) & (((potted >> np.arange(len(frame_types))) & 1) == 0)).astype(np.uint8)
            rest_pos = np.where(present[:, None].astype(bool), fpos, 0.0).astype(np.float32)
            sol = dict(cost=int(cost), breaks=level + 1, strikes=[(p[2], p[3]) for p in path], dirs=[p[0] for p in path],
                       speeds=[p[1] for p in path], frames=[(p[4], p[5]) for p in path], pos=rest_pos, present=present)
            if level + 1 >= levels:                                         # simCount == BREAKS (:1407)
                solutions.append(sol)
                continue
            search = rules.next_search(frame_types, present, parent["searched"])
            ok, pos2, present2 = rules.state_after_break(rest_pos, frame_types, present)
            if not ok:                                                       # "could not replace all balls" (:1432-1435)
                continue
            on = present2.astype(bool)
            if any(on[i] and int(frame_types[i]) in search for i in range(len(frame_types))):   # hasTypeToSearch (:1437-1443)
                nxt.append(dict(pos=np.where(on[:, None], pos2, 0.0).astype(np.float32),
                                types=np.where(on, frame_types, -1).astype(np.int32), searched=search, cost=int(cost),
                                path=path))
            else:
                sol["pos"], sol["present"] = pos2, present2
                solutions.append(sol)
        states = nxt
    solutions.sort(key=lambda s: s["cost"])   # stable: equal costs stay in discovery order
    return solutions


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
            key = id(fpos)
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
