"""CPU tier, world size 2 over gloo: the host-side logic of the multi-GPU path -- contiguous candidate
shards, per-rank top-k by (cost, id), all-gather of the record blocks, identical merge on every rank -- against a
single-process answer.  The device kernels are not involved (no GPU here); the record-block layout and the merge
rule are the ones of billiard_ai_b200/csrc/bcu_merge.cu, restated in numpy by tests/merge_restatement.py."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import merge_restatement as D
from billiard_ai_b200 import workloads as W


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _fake_level(total, nb, seed=4):
    """Synthetic per-candidate results with many cost ties, so that the id tie-break is exercised."""
    rng = np.random.RandomState(seed)
    cost = rng.randint(0, 50, total).astype(np.uint64) * 1000
    valid = rng.rand(total) < 0.3
    potted = rng.randint(0, 1 << nb, total).astype(np.uint32)
    time = rng.rand(total).astype(np.float32)
    pos = rng.rand(total, nb, 2).astype(np.float32)
    return cost, valid, potted, time, pos


def _worker(rank, world, port, total, nb, k, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    cost, valid, potted, time, pos = _fake_level(total, nb)
    lo, hi = W.shard_range(total, rank, world)
    block = D.pack_local_topk(cost[lo:hi], valid[lo:hi], potted[lo:hi], time[lo:hi], pos[lo:hi], first_id=lo, k=k)
    gathered = [torch.zeros_like(torch.from_numpy(block)) for _ in range(world)]
    dist.all_gather(gathered, torch.from_numpy(block))
    merged = D.merge_blocks([g.numpy() for g in gathered], k, nb)
    np.save(os.path.join(out_dir, f"merged_{rank}.npy"), merged)
    dist.destroy_process_group()


def test_two_rank_shard_allgather_merge_equals_single_process(tmp_path):
    total, nb, k, world = 5000, 6, 64, 2
    mp.spawn(_worker, args=(world, _free_port(), total, nb, k, str(tmp_path)), nprocs=world, join=True)
    m0 = np.load(tmp_path / "merged_0.npy")
    m1 = np.load(tmp_path / "merged_1.npy")
    assert np.array_equal(m0, m1)                      # every rank holds the identical beam: no broadcast needed
    cost, valid, potted, time, pos = _fake_level(total, nb)
    single = D.pack_local_topk(cost, valid, potted, time, pos, first_id=0, k=k)
    assert np.array_equal(m0, single)                  # == the single-GPU answer
    v = D.view_block(m0, k, nb)
    ids = np.nonzero(valid)[0]
    order = np.lexsort((ids, cost[ids]))[:k]
    assert np.array_equal(v["id"], ids[order].astype(np.uint64))
    assert np.array_equal(v["pos"], pos[ids[order]])


def test_merge_handles_short_and_empty_ranks():
    nb, k = 4, 16
    cost, valid, potted, time, pos = _fake_level(40, nb, seed=9)
    valid[:] = False
    valid[[3, 17, 35]] = True
    blocks = [D.pack_local_topk(cost[lo:hi], valid[lo:hi], potted[lo:hi], time[lo:hi], pos[lo:hi], first_id=lo, k=k)
              for lo, hi in (W.shard_range(40, r, 4) for r in range(4))]
    v = D.view_block(D.merge_blocks(blocks, k, nb), k, nb)
    n = int((v["id"] != np.uint64(2**64 - 1)).sum())
    assert n == 3 and set(v["id"][:3].tolist()) == {3, 17, 35}
    assert np.all(np.diff(v["cost"][:3].astype(np.int64)) >= 0)
