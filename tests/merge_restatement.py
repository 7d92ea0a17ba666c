"""TEST INFRASTRUCTURE: a numpy restatement of the multi-GPU level merge -- the checker of the device
merge kernel (tests/test_gpu_full_size.py, through bcu_test_merge_records) and of the sharding rule under gloo
(tests/test_distributed_cpu.py).  Not part of the product.

The product path does this on the device (csrc/bcu_topk.cu, csrc/bcu_merge.cu: top-k -> record block ->
ncclAllGather -> merge).  This module restates the record-block layout and the merge rule on the host so that the
sharding / all-gather / merge logic can be exercised with a world-size-2 gloo group on CPU (tests/test_distributed_cpu.py)
and so that host callers can decode a block fetched with bcu_fetch_beam-style layouts.

Record block of k entries, N balls (little endian, contiguous):
    cost u64[k] | id u64[k] | potted u32[k] | time f32[k] | pos f32[k][N][2]        entries past the valid count: cost = id = ~0
Order: ascending (cost, id) -- the reference's solution order (process.hpp:19-25) with the global candidate id as tie-break.
"""
import numpy as np

NONE = np.uint64(2**64 - 1)


def block_bytes(k, n_balls):
    return k * (8 + 8 + 4 + 4 + n_balls * 8)


def view_block(block, k, n_balls):
    b = np.ascontiguousarray(block, dtype=np.uint8)
    assert b.size == block_bytes(k, n_balls)
    o = 0
    out = {}
    for name, dt, cnt in (("cost", np.uint64, k), ("id", np.uint64, k), ("potted", np.uint32, k), ("time", np.float32, k),
                          ("pos", np.float32, k * n_balls * 2)):
        nbytes = np.dtype(dt).itemsize * cnt
        out[name] = b[o:o + nbytes].view(dt)
        o += nbytes
    out["pos"] = out["pos"].reshape(k, n_balls, 2)
    return out


def pack_local_topk(cost, valid, potted, time, pos, first_id, k):
    """Local top-k of one rank's shard -> record block (what topk_pass + pack_records produce on the device)."""
    n_balls = pos.shape[1]
    block = np.zeros(block_bytes(k, n_balls), np.uint8)
    v = view_block(block, k, n_balls)
    v["cost"][:] = NONE
    v["id"][:] = NONE
    idx = np.nonzero(valid)[0]
    order = idx[np.lexsort((idx, cost[idx]))][:k]
    m = len(order)
    v["cost"][:m] = cost[order]
    v["id"][:m] = order.astype(np.uint64) + np.uint64(first_id)
    v["potted"][:m] = potted[order]
    v["time"][:m] = time[order]
    v["pos"][:m] = pos[order]
    return block


def merge_blocks(blocks, k, n_balls):
    """G sorted blocks -> the global top-k block (merge_records on the device): deterministic, identical on every rank."""
    views = [view_block(b, k, n_balls) for b in blocks]
    cost = np.concatenate([v["cost"] for v in views])
    ids = np.concatenate([v["id"] for v in views])
    keep = np.nonzero(ids != NONE)[0]
    order = keep[np.lexsort((ids[keep], cost[keep]))][:k]
    out = np.zeros(block_bytes(k, n_balls), np.uint8)
    o = view_block(out, k, n_balls)
    o["cost"][:] = NONE
    o["id"][:] = NONE
    m = len(order)
    o["cost"][:m] = cost[order]
    o["id"][:m] = ids[order]
    o["potted"][:m] = np.concatenate([v["potted"] for v in views])[order]
    o["time"][:m] = np.concatenate([v["time"] for v in views])[order]
    o["pos"][:m] = np.concatenate([v["pos"] for v in views])[order]
    return out
