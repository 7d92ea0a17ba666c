"""Comparison of engine results with oracle results under the north_star parity rules:

  * the event sequence (event type, ball and rail/pocket ids, order) must be identical up to the first event
    whose time is within 1e-9 s of a competing event (near tie: `min_gap` from the port oracle) -- such
    candidates are excluded from the strict comparison;
  * positions / times must agree within 1e-6 relative (rtol below);  costs, flags, potted masks exactly.

TEST INFRASTRUCTURE ONLY.
"""
import numpy as np

RTOL = 1e-6      # north_star: positions and velocities within 1e-6 relative
TIE_EPS = 1e-9   # north_star: tie clause


def rel_close(a, b, rtol=RTOL, scale=None):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    s = np.maximum(np.abs(a), np.abs(b)) if scale is None else scale
    return np.abs(a - b) <= rtol * np.maximum(s, 1e-30)


def compare_batch(gpu, ora, table_scale=1881.0):
    """gpu / ora: dicts with cost, flags, n_events, potted, final_time, final_pos, event_log (optional), ora may
    have min_gap.  Returns dict of boolean arrays (per candidate) and summary numbers."""
    n = len(gpu["cost"])
    clean = np.ones(n, bool)
    if ora.get("min_gap") is not None:
        clean = ora["min_gap"] >= TIE_EPS
    same_struct = (gpu["flags"] == ora["flags"]) & (gpu["n_events"] == ora["n_events"]) & (gpu["potted"] == ora["potted"])
    same_cost = gpu["cost"] == ora["cost"]
    # positions are coordinates on a 1881 x 943 mm table centred at 0: relative to the table scale
    pos_ok = np.ones(n, bool)
    if gpu.get("final_pos") is not None and ora.get("final_pos") is not None:
        pos_ok = rel_close(gpu["final_pos"], ora["final_pos"], scale=table_scale).reshape(n, -1).all(axis=1)
    time_ok = rel_close(gpu["final_time"], ora["final_time"])
    bits_ok = np.ones(n, bool)
    if gpu.get("final_pos") is not None and ora.get("final_pos") is not None:
        bits_ok = (gpu["final_pos"].view(np.uint32) == ora["final_pos"].view(np.uint32)).reshape(n, -1).all(axis=1)
        bits_ok &= gpu["final_time"].view(np.uint32) == ora["final_time"].view(np.uint32)
    log_ok = np.ones(n, bool)
    if gpu.get("event_log") is not None and ora.get("event_log") is not None:
        a, b = gpu["event_log"], ora["event_log"]
        coll = b[..., 0] == 0
        lo = np.where(coll, np.minimum(b[..., 1], b[..., 2]), b[..., 1])
        hi = np.where(coll, np.maximum(b[..., 1], b[..., 2]), b[..., 2])
        same = (a[..., 0] == b[..., 0]) & (a[..., 1] == lo) & (a[..., 2] == hi)
        ta = a[..., 3].copy().view(np.float32)
        tb = b[..., 3].copy().view(np.float32)
        valid = b[..., 0] >= 0
        tok = rel_close(np.where(valid, ta, 0), np.where(valid, tb, 0))
        log_ok = (same & tok).all(axis=1)
    ok = same_struct & same_cost & pos_ok & time_ok & log_ok
    return dict(clean=clean, ok=ok, same_struct=same_struct, same_cost=same_cost, pos_ok=pos_ok, time_ok=time_ok,
                log_ok=log_ok, bit_exact=ok & bits_ok,
                n=n, n_clean=int(clean.sum()), n_ok=int((ok & clean).sum()), n_bit=int((ok & bits_ok & clean).sum()))


def summary(c):
    nc = max(c["n_clean"], 1)
    return (f"{c['n']} candidates, {c['n_clean']} clear of near-ties; parity {c['n_ok']}/{c['n_clean']} "
            f"({100.0 * c['n_ok'] / nc:.3f}%), bit-exact {c['n_bit']}/{c['n_clean']} ({100.0 * c['n_bit'] / nc:.3f}%)")
