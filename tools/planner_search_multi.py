#!/usr/bin/env python
"""bcu_search_levels in planner mode on N GPUs (torchrun): every rank plans the whole beam (the plan is deterministic), simulates
its contiguous share, the per-rank beams are merged by the engine's one allgather per level.  Checks: all ranks hold the same
solution list; Situation 2 with the snooker rules gives the thesis' two-break totals; a 16-ball table gives the same digest
for every N (printed).

    python tools/planner_search_multi.py
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/planner_search_multi.py"""
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch   # noqa: E402
import torch.distributed as dist   # noqa: E402
from billiard_ai_b200 import Engine, workloads as W   # noqa: E402
from billiard_ai_b200.rules import RepeatRules, SnookerRules   # noqa: E402
from billiard_ai_b200.search import search_levels   # noqa: E402


def main():
    rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    tbl = W.default_table()
    eng = Engine.from_table(tbl, W.SNOOKER_SCORES, cue_type=W.TYPE_CODE["WHITE"], device=local)
    if world > 1:
        box = [eng.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        eng.comm_init(world, rank, box[0])
    eng.set_planner(tbl, velocity_multiply=False)
    with open(os.path.join(ROOT, "tests", "golden", "layouts.json")) as f:
        lay = json.load(f)["situation2"]
    with open(os.path.join(ROOT, "tests", "golden", "thesis_costs.json")) as f:
        thesis = json.load(f)["lists"]["s2_two_breaks"]
    pos = np.array([b["position"] for b in lay["balls"]], np.float32)
    types = np.array([W.TYPE_CODE[b["type"]] for b in lay["balls"]], np.int32)
    sols = search_levels(eng, pos, types, [W.TYPE_CODE["RED"]], 2, 1024, None, None, SnookerRules())
    costs = [s["cost"] for s in sols]
    assert costs[:len(thesis)] == thesis, (costs[:8], thesis)
    eng.set_planner(tbl, velocity_multiply=True)
    p16, t16 = W.random_layout(16, 16)
    t0 = time.perf_counter()
    stats, info = [], {}
    big = search_levels(eng, p16, t16, [W.TYPE_CODE["RED"]], 3, 1024, None, None, RepeatRules(), stats=stats, info=info)
    wall = time.perf_counter() - t0
    digest = hashlib.sha256(json.dumps([(s["cost"], s["strikes"]) for s in big]).encode()).hexdigest()[:16]
    if world > 1:
        allv = [None] * world
        dist.all_gather_object(allv, (digest, costs[:8]))
        assert len(set(json.dumps(v) for v in allv)) == 1, "ranks disagree"
    if rank == 0:
        print(json.dumps({"n_gpus": world, "situation2_two_breaks": costs[:8], "thesis": thesis, "pool16_three_breaks_planned": {
            "solutions": len(big), "best": [s["cost"] for s in big[:4]], "digest": digest, "levels": stats, "wall_s": wall, "info": info}}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
