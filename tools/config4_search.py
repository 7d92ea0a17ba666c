#!/usr/bin/env python
"""BASELINE config 4: snooker 22-ball frame, 4-break search, ~1 M candidate strikes per level (beam 1024 x 32 x 32),
sharded over the ranks of a torchrun launch with the engine's one ncclAllGather per level.

    python tools/config4_search.py                      # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/config4_search.py

Prints per-level device times and the best solutions on rank 0; every rank asserts that it holds the same solution list."""
import hashlib, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from billiard_ai_b200 import Engine, workloads as W
from billiard_ai_b200.rules import SnookerRules
from billiard_ai_b200.search import search_levels


def main():
    rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = Engine.from_table(W.default_table(), W.SNOOKER_SCORES, cue_type=W.TYPE_CODE["WHITE"], device=local)
    if world > 1:
        box = [eng.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        eng.comm_init(world, rank, box[0])
    pos, types = W.snooker22_open()
    dirs, speeds = W.grid(32, 32)
    levels, beam = 4, 1024
    for it in range(2):   # first pass warms up (allocations, graph capture)
        stats = []
        t0 = time.perf_counter()
        sols = search_levels(eng, pos, types, [W.TYPE_CODE["RED"]], levels, beam, dirs, speeds, SnookerRules(), rank=rank, world=world, stats=stats)
        wall = time.perf_counter() - t0
    digest = hashlib.sha256(json.dumps([(s["cost"], s["dirs"], s["speeds"]) for s in sols]).encode()).hexdigest()
    if world > 1:
        allv = [None] * world
        dist.all_gather_object(allv, digest)
        assert len(set(allv)) == 1, "ranks disagree on the solution list"
    if rank == 0:
        strikes = sum(s["candidates"] for s in stats)
        print(json.dumps({"config": "config4: snooker 22-ball, 4 breaks, beam 1024 x 32 dirs x 32 speeds", "n_gpus": world,
                          "levels": stats, "strikes": strikes, "wall_s": wall, "strikes_per_s_wall": strikes / wall,
                          "solutions": len(sols), "best": [(s["cost"], s["breaks"]) for s in sols[:3]], "digest": digest[:16]}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
