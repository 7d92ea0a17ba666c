"""bcu_topk_static (include/billiard_cuda.h): the top-k selection for callers that cannot allocate -- scratch in a file-scope
__device__ array, inputs and outputs as device pointers.  No counterpart in the reference;
checked here, through the C ABI, against the existing entry point bcu_topk on the same launch results -- bit for
bit, the bar of every selection test in this repository -- on two calls in a row, then on the dense path of the selection
and on the size limits.

Collected last on purpose.  Device memory for the caller's side comes from torch, which is
what it is here for; the entry point itself is given raw pointers and a stream.
"""
import ctypes as C

import numpy as np
import pytest

from oracle_lib import TYPE_CODE, SNOOKER_SCORES, F_VALID

pytestmark = pytest.mark.gpu
PAD = np.uint64(2**64 - 1)
_calls = 0


@pytest.fixture(scope="module")
def engine(table):
    from billiard_ai_b200 import Engine
    e = Engine.from_table(table, SNOOKER_SCORES, cue_type=TYPE_CODE["WHITE"])
    yield e
    e.close()


def _static_topk(cost, flags, k, first=0):
    """Uploads (cost, flags), runs bcu_topk_static on torch's current stream and returns (costs [k], ids [k], valid count)."""
    import torch
    from billiard_ai_b200.engine import topk_static
    global _calls
    dev = torch.device("cuda:0")
    n = len(cost)
    d_cost = torch.zeros(max(n, 1), dtype=torch.int64, device=dev)      # never a null pointer, also for n = 0
    d_flags = torch.zeros(max(n, 1), dtype=torch.int32, device=dev)
    d_cost[:n] = torch.from_numpy(cost.view(np.int64)).to(dev)
    d_flags[:n] = torch.from_numpy(flags.view(np.int32)).to(dev)
    # outputs start as garbage on purpose: the call must write all k entries and the count
    d_oc = torch.full((k,), 0x5a5a5a5a, dtype=torch.int64, device=dev)
    d_oi = torch.full((k,), 0x5a5a5a5a, dtype=torch.int64, device=dev)
    d_n = torch.full((1,), -77, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    torch.cuda.synchronize(dev)
    free_before = torch.cuda.mem_get_info(dev)[0]
    topk_static(d_cost.data_ptr(), d_flags.data_ptr(), n, first, k, d_oc.data_ptr(), d_oi.data_ptr(), d_n.data_ptr(), stream)
    torch.cuda.synchronize(dev)
    # no device memory is taken by the call (the first call of a process may load the module that holds the kernels and the
    # static array -- CUDA loads modules lazily -- so the figure is checked from the second call on)
    if _calls > 0:
        assert torch.cuda.mem_get_info(dev)[0] >= free_before
    _calls += 1
    return d_oc.cpu().numpy().view(np.uint64), d_oi.cpu().numpy().view(np.uint64), int(d_n.item())


def _launch(engine, n_balls, seed, n_dirs, n_speeds):
    from billiard_ai_b200 import workloads as W
    pos, types = W.random_layout(n_balls, seed)
    dirs, speeds = W.grid(n_dirs, n_speeds)
    engine.set_states(pos, types, 1 << TYPE_CODE["RED"])
    engine.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
    return engine.fetch(final_pos=False)


def _check_against_existing(engine, res, k):
    want_cost, want_id = engine.topk(k)          # the existing entry point, on the same launch
    got_cost, got_id, n_valid = _static_topk(np.ascontiguousarray(res["cost"]), np.ascontiguousarray(res["flags"]), k)
    assert n_valid == int(((res["flags"] & F_VALID) != 0).sum())
    m = len(want_cost)
    assert m == min(k, n_valid)
    assert np.array_equal(got_cost[:m], want_cost) and np.array_equal(got_id[:m], want_id)
    assert np.all(got_cost[m:] == PAD) and np.all(got_id[m:] == PAD)


def test_static_topk_matches_bcu_topk_on_two_calls_in_a_row(engine):
    """Two different launches, two different k, one after the other on the shared scratch: the second call must not see
    anything of the first (and a third, repeating the first, nothing of the second)."""
    res_a = _launch(engine, 16, 16, 256, 64)
    assert int(((res_a["flags"] & F_VALID) != 0).sum()) > 100
    _check_against_existing(engine, res_a, 1024)
    res_b = _launch(engine, 12, 7, 256, 32)
    _check_against_existing(engine, res_b, 100)
    res_a2 = _launch(engine, 16, 16, 256, 64)
    _check_against_existing(engine, res_a2, 1024)
    _check_against_existing(engine, res_a2, 7)


def test_static_topk_dense_path_then_sparse_again():
    """More valid keys than the compacted list holds (the streaming sorter and the rank merge, whose per-CTA lists are the
    larger part of the static scratch), against numpy's sort of (cost, id); ties in cost are broken by id.  Then a sparse
    call right behind it."""
    rng = np.random.default_rng(5)
    for n, share, k, first in ((300_000, 0.6, 1024, 0), (1 << 20, 0.9, 513, 1 << 33), (50_000, 0.02, 64, 12345), (300_000, 0.6, 1, 0)):
        cost = rng.integers(0, 5000, size=n, dtype=np.uint64)      # heavy ties
        flags = np.where(rng.random(n) < share, 3, 1).astype(np.uint32) | (rng.integers(0, 31, size=n).astype(np.uint32) << 8)
        got_cost, got_id, n_valid = _static_topk(cost, flags, k, first)
        idx = np.nonzero((flags & F_VALID) != 0)[0]
        order = np.lexsort((idx, cost[idx]))[:k]
        assert n_valid == len(idx)
        m = min(k, len(idx))
        assert np.array_equal(got_cost[:m], cost[idx][order]) and np.array_equal(got_id[:m], (idx[order] + first).astype(np.uint64))
        assert np.all(got_cost[m:] == PAD) and np.all(got_id[m:] == PAD)


def test_static_topk_refuses_what_the_scratch_cannot_serve():
    import torch
    from billiard_ai_b200 import BcuError
    from billiard_ai_b200.engine import topk_static
    dev = torch.device("cuda:0")
    t = torch.zeros(16, dtype=torch.int64, device=dev)
    p = t.data_ptr()
    with pytest.raises(BcuError, match="beyond the maximum.*4194305 candidates"):
        topk_static(p, p, (1 << 22) + 1, 0, 8, p, p, p)
    with pytest.raises(BcuError, match="beyond the maximum.*k = 1025"):
        topk_static(p, p, 16, 0, 1025, p, p, p)
    with pytest.raises(BcuError, match="null pointer"):
        topk_static(p, 0, 16, 0, 8, p, p, p)
    # the limits themselves are served, and an empty input gives an empty, padded answer
    cost = np.arange(1 << 22, dtype=np.uint64)[::-1].copy()
    flags = np.full(1 << 22, 3, np.uint32)
    got_cost, got_id, n_valid = _static_topk(cost, flags, 1024)
    assert n_valid == 1 << 22 and np.array_equal(got_cost, np.arange(1024, dtype=np.uint64))
    assert np.array_equal(got_id, (1 << 22) - 1 - np.arange(1024, dtype=np.uint64))
    got_cost, got_id, n_valid = _static_topk(np.zeros(0, np.uint64), np.zeros(0, np.uint32), 5)
    assert n_valid == 0 and np.all(got_cost == PAD) and np.all(got_id == PAD)
