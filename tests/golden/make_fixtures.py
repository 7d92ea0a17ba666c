#!/usr/bin/env python
"""Generates the committed golden fixtures under tests/golden/ from the reference tree.

Run in the build container only (needs /root/reference and oracle/_ref/libref_oracle.so):
    python tests/golden/make_fixtures.py

Outputs (all small, committed):
  table.json          table geometry of test/billiard_search/resources/configuration.json (+ the rail
                      normals / shifted end points exactly as the reference's Rail ctor computes them)
  layouts.json        resources/00_test_01.json state; thesis "Situation 1/2" states
  filmed_shots.json   the filmed-shot fixtures of test/billiard_search/src/simulation_vs_reality.cpp:
                      ball positions, cue velocity (calculateInitialVelocity), asserted layer count and
                      asserted node type / isRolling per layer
  thesis_costs.json   the [mapToSolution] cost logs in doc_source/.../04_suche.tex
  ref_vectors.npz     outputs of the reference itself (libref_oracle.so) on seeded random inputs
"""
import json
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
REF = "/root/reference"


def f32(x):
    return float(np.float32(x))


def make_table():
    with open(f"{REF}/test/billiard_search/resources/configuration.json") as f:
        cfg = json.load(f)
    tbl = {
        "source": "reference test/billiard_search/resources/configuration.json",
        "ballRadius": cfg["ball"]["radius"],
        "minimalPocketVelocity": cfg["table"]["minimalPocketVelocity"],
        "innerTableLength": cfg["table"]["innerTableLength"],
        "innerTableWidth": cfg["table"]["innerTableWidth"],
        "rails": [{"id": r["id"], "start": [r["start"]["x"], r["start"]["y"]], "end": [r["end"]["x"], r["end"]["y"]],
                   "shift": [r["shiftDirection"]["x"], r["shiftDirection"]["y"]]} for r in cfg["banks"]],
        "pockets": [{"id": p["id"], "position": [p["position"]["x"], p["position"]["y"]], "radius": p["radius"],
                     "normal": [p["normal"]["x"], p["normal"]["y"]], "type": p["type"],
                     "pottingPoint": [p["pottingPoints"][0]["x"], p["pottingPoints"][0]["y"]]} for p in cfg["pockets"]],
    }
    return tbl


def parse_vec_defs(block):
    """glm::vec2 definitions inside one TEST body -> dict name -> (x, y) evaluated in float32."""
    env = {}
    pat = re.compile(r"glm::vec2&?\s+(\w+)\s*=\s*(.+?);")
    for m in pat.finditer(block):
        name, expr = m.group(1), m.group(2).strip()
        lit = re.match(r"glm::vec2\s*\{\s*([-\d.eE]+)\s*,\s*([-\d.eE]+)\s*\}", expr)
        if lit:
            env[name] = (np.float32(float(lit.group(1))), np.float32(float(lit.group(2))))
            continue
        sub = re.match(r"(\w+)\s*-\s*(\w+)$", expr)
        if sub and sub.group(1) in env and sub.group(2) in env:
            a, b = env[sub.group(1)], env[sub.group(2)]
            env[name] = (np.float32(a[0] - b[0]), np.float32(a[1] - b[1]))
    return env


NODE_KIND = {"BallMovingNode": 0, "BallCollisionNode": 1, "BallRailCollisionNode": 2, "BallPottingNode": 3,
             "BallShotNode": 4, "BallInRestNode": 5}


def parse_filmed_shots(ref):
    src = open(f"{REF}/test/billiard_search/src/simulation_vs_reality.cpp").read()
    tests = []
    for m in re.finditer(r"^TEST\((\w+),\s*(\w+)\)\s*\{", src, re.M):
        suite, name = m.group(1), m.group(2)
        start = m.end()
        nxt = re.search(r"^TEST\(|^std::ostream& operator<<", src[start:], re.M)
        block = src[start:start + nxt.start()] if nxt else src[start:]
        env = parse_vec_defs(block)
        vm = re.search(r"glm::vec2 velocity = calculateInitialVelocity\((\w+)\s*-\s*(\w+),\s*(\d+)\)", block)
        if not vm:
            continue
        a, b = env[vm.group(1)], env[vm.group(2)]
        dx, dy = np.float32(a[0] - b[0]), np.float32(a[1] - b[1])
        vx, vy = ref.initial_velocity(float(dx), float(dy), int(vm.group(3)))
        ids = dict(re.findall(r'std::string (\w+) = "([\w-]+)";', block))
        balls = []
        for bm in re.finditer(r'billiard::search::Ball\s*\{\s*(\w+),\s*"(\w+)",\s*(\w+)\s*\}', block):
            p = env[bm.group(1)]
            balls.append({"var": bm.group(3), "id": ids.get(bm.group(3), bm.group(3)), "type": bm.group(2),
                          "position": [float(p[0]), float(p[1])]})
        layers = [int(x) for x in re.findall(r"EXPECT_EQ\(system\._layers\.size\(\), (\d+)\)", block)]
        alt = re.findall(r"system\._layers\.size\(\) == (\d+)", block)
        layers += [int(x) for x in alt]
        layer_vars = {lm.group(1): int(lm.group(2))
                      for lm in re.finditer(r"Layer&\s+(\w+)\s*=\s*system\._layers\[(\d+)\];", block)}
        expects = []
        node_vars = {}
        # `if (system._layers.size() == K) { ... } else { ... }` regions make expectations conditional
        regions = []
        for im in re.finditer(r"if \(system\._layers\.size\(\) == (\d+)\) \{", block):
            k = int(im.group(1))
            els = block.find("} else {", im.end())
            end = block.find("\n    }", els) if els >= 0 else -1
            regions.append((im.end(), els, end, k))

        def condition(posn):
            for a, els, end, k in regions:
                if a <= posn < els:
                    return {"onlyIfLayers": k}
                if els >= 0 and els <= posn < end:
                    return {"notIfLayers": k}
            return {}

        for nm in re.finditer(r"auto (\w+) = getNodeOfBall<(\w+), &Node::\w+>\((\w+), (\w+)\);", block):
            var, kind, layer_var, ball_var = nm.groups()
            if layer_var not in layer_vars:
                continue
            asserted = re.search(r"ASSERT_TRUE\(%s\);" % var, block) is not None
            node_vars[var] = (layer_vars[layer_var], ball_var, kind)
            if asserted:
                expects.append(dict({"layer": layer_vars[layer_var], "ball": ball_var, "kind": NODE_KIND[kind]},
                                    **condition(nm.start())))
        for rm in re.finditer(r"EXPECT_EQ\((\w+)->_after\._isRolling, (true|false)\);", block):
            if rm.group(1) in node_vars:
                layer, ball_var, kind = node_vars[rm.group(1)]
                expects.append({"layer": layer, "ball": ball_var, "kind": NODE_KIND[kind],
                                "afterRolling": rm.group(2) == "true"})
        tests.append({"name": name, "enabled": not suite.startswith("DISABLED_"), "balls": balls,
                      "velocity": [vx, vy], "frames": int(vm.group(3)), "expectedLayers": sorted(set(layers)),
                      "expects": expects})
    return tests


def parse_thesis():
    tex = open(f"{REF}/doc_source/tex/bachelor_thesis/de/04_results/04_suche.tex").read().splitlines()
    logs = []
    cur = None
    for i, line in enumerate(tex, 1):
        if "[mapToSolution] COST START" in line:
            cur = {"line": i, "search": [], "simulation": []}
        elif cur is not None and "cost-search" in line:
            m = re.search(r"cost: (\d+) previous cost: (\d+) total cost: (\d+)", line)
            cur["search"].append([int(m.group(1)), int(m.group(3))])
        elif cur is not None and "cost-simulation" in line:
            m = re.search(r"cost: (\d+) previous cost: (\d+)", line)  # some logs are interleaved with other threads
            cur["simulation"].append([int(m.group(1)), int(m.group(2))])
        elif cur is not None and "end cost" in line:
            cur["end"] = int(re.search(r"end cost: (\d+)", line).group(1))
            logs.append(cur)
            cur = None

    def states(start, end):
        out = []
        for line in tex[start - 1:end]:
            m = re.match(r"%(\w+), (\w+), ([-\d.]+), ([-\d.]+)", line)
            if m:
                out.append({"id": m.group(1), "type": m.group(2), "position": [float(m.group(3)), float(m.group(4))]})
        return out

    return {
        "source": "doc_source/tex/bachelor_thesis/de/04_results/04_suche.tex",
        "situation1": states(27, 33), "situation2": states(986, 992),
        "lists": {
            "s1_single_no_banks": [l["end"] for l in logs if 100 < l["line"] < 300],
            "s1_three_speeds": [l["end"] for l in logs if 400 < l["line"] < 620],
            "s1_with_banks": [l["end"] for l in logs if 700 < l["line"] < 960],
            "s2_two_breaks": [l["end"] for l in logs if 1000 < l["line"] < 1300],
        },
        "logs": logs,
    }


def main():
    from oracle_lib import RefOracle, TYPE_CODE

    tbl = make_table()
    ref = RefOracle(tbl)
    tbl["railsDerived"] = [[float(x) for x in row] for row in ref.rail_derived()]
    tbl["railsDerivedLayout"] = "normal.xy, shiftedStart.xy, shiftedEnd.xy (reference type.cpp:56-72)"
    with open(os.path.join(HERE, "table.json"), "w") as f:
        json.dump(tbl, f, indent=1)

    with open(f"{REF}/test/billiard_search/resources/00_test_01.json") as f:
        t01 = json.load(f)
    thesis = parse_thesis()
    layouts = {
        "test_01": {"source": "test/billiard_search/resources/00_test_01.json",
                    "balls": [{"id": b["id"], "type": b["type"], "position": [b["position"]["x"], b["position"]["y"]]}
                              for b in t01["state"]["balls"]], "searchTypes": [t01["search"]["type"]]},
        "situation1": {"source": thesis["source"] + ":27-33", "balls": thesis["situation1"], "searchTypes": ["RED"]},
        "situation2": {"source": thesis["source"] + ":986-992", "balls": thesis["situation2"], "searchTypes": ["RED"]},
    }
    with open(os.path.join(HERE, "layouts.json"), "w") as f:
        json.dump(layouts, f, indent=1)
    with open(os.path.join(HERE, "thesis_costs.json"), "w") as f:
        json.dump({k: v for k, v in thesis.items() if k not in ("situation1", "situation2")}, f, indent=1)

    shots = parse_filmed_shots(ref)
    with open(os.path.join(HERE, "filmed_shots.json"), "w") as f:
        json.dump({"source": "test/billiard_search/src/simulation_vs_reality.cpp", "tests": shots}, f, indent=1)
    print("filmed shots:", len(shots), "enabled:", sum(t["enabled"] for t in shots))

    # Reference planner candidates + results for the published layouts (config 1 and thesis lists).
    plans = {}
    for key, lay in layouts.items():
        pos = np.array([b["position"] for b in lay["balls"]], dtype=np.float32)
        types = np.array([TYPE_CODE[b["type"]] for b in lay["balls"]], dtype=np.int32)
        mask = sum(1 << TYPE_CODE[t] for t in lay["searchTypes"])
        for vm in (0, 1):
            ref.set_toggles(0, vm)
            pl = ref.plan(pos, types, mask)
            plans[f"{key}_vm{vm}"] = pl
            sols = np.sort(pl["result_cost"][pl["result_cost"] != np.uint64(2**64 - 1)])
            print(key, "velocityMultiply", vm, "simulations", len(pl["vx"]), "solutions", len(sols), sols[:10])
    ref.set_toggles(0, 1)
    flat = {}
    for k, pl in plans.items():
        for kk, vv in pl.items():
            flat[f"plan_{k}_{kk}"] = np.asarray(vv)

    # Reference outputs on seeded random inputs (16-ball layouts, random cue velocities).
    rng = np.random.RandomState(20261104)
    n = 16
    pos = []
    while len(pos) < n:
        p = np.array([rng.uniform(-880, 880), rng.uniform(-420, 420)], dtype=np.float32)
        if all(np.hypot(*(p - q)) > 53.3 for q in pos):
            pos.append(p)
    pos = np.array(pos, dtype=np.float32)
    types = np.array([0] + [1] * 9 + [2, 3, 4, 5, 6, 7], dtype=np.int32)
    nc = 512
    ang = rng.uniform(0, 2 * np.pi, nc)
    spd = rng.uniform(300, 5000, nc)
    vx = (spd * np.cos(ang)).astype(np.float32)
    vy = (spd * np.sin(ang)).astype(np.float32)
    ev = ref.evaluate(pos[None], types[None], 1 << 1, 0, vx, vy, 100000, 0, threads=8, event_log=True)
    flat.update(rand16_pos=pos, rand16_types=types, rand16_vx=vx, rand16_vy=vy, rand16_cost=ev["cost"],
                rand16_flags=ev["flags"], rand16_n_events=ev["n_events"], rand16_potted=ev["potted"],
                rand16_final_time=ev["final_time"], rand16_final_pos=ev["final_pos"], rand16_event_log=ev["event_log"])
    np.savez_compressed(os.path.join(HERE, "ref_vectors.npz"), **flat)
    print("rand16: completed", int((ev["flags"] & 1).sum()), "valid", int(((ev["flags"] & 2) != 0).sum()),
          "mean events", ev["n_events"].mean())


if __name__ == "__main__":
    main()
