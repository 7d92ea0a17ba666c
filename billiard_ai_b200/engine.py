"""ctypes binding of libbilliard_cuda.so (include/billiard_cuda.h).

Host-side mirror of the reference's evaluation entry points for this path:
  Engine.simulate_trace  <->  billiard::search::simulate(State, vec2, Configuration)   billiard_search.cpp:141-175
  Engine.evaluate        <->  one SIMULATION node per candidate: simulate(node, state)  billiard_search.cpp:1358-1471
                              + simulationCost (:1473-1532) + rules validEndState (billiard_snooker.cpp:1128-1153)
  Engine.evaluate_grid / topk / gather  <->  one break level of the search (expand + best-first pop order,
                              process.hpp:19-25) over a synthetic direction x speed grid

There is no CPU path: a missing library or a missing CUDA device raises BcuError.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BCU_LIB") or os.path.join(HERE, "libbilliard_cuda.so")   # BCU_LIB: tuning builds (tools/build_variant.sh)

F_COMPLETED, F_VALID, F_OVERFLOW, F_NOEVENT = 1, 2, 4, 8
MAX_LAYERS = 32
TRACE_REC = 10

_STATUS = {1: "BCU_ERR_NO_DEVICE", 2: "BCU_ERR_INVALID", 3: "BCU_ERR_CUDA", 4: "BCU_ERR_STATE"}


class BcuError(RuntimeError):
    pass


class _TableDesc(C.Structure):
    _fields_ = [("n_rails", C.c_int32), ("n_pockets", C.c_int32), ("ball_radius", C.c_float),
                ("rails", C.POINTER(C.c_float)), ("pockets", C.POINTER(C.c_float)),
                ("rails_derived", C.POINTER(C.c_float))]


class _RulesDesc(C.Structure):
    _fields_ = [("n_types", C.c_int32), ("cue_type", C.c_int32), ("score", C.POINTER(C.c_double))]


class _PlannerDesc(C.Structure):
    _fields_ = [("n_rails", C.c_int32), ("rail_start", C.POINTER(C.c_float)), ("rail_order", C.POINTER(C.c_int32)),
                ("n_set_order", C.c_int32), ("set_order", C.POINTER(C.c_int32)), ("n_search_pockets", C.c_int32),
                ("search_pockets", C.POINTER(C.c_float)), ("minimal_pocket_velocity", C.c_float),
                ("diagonal_length_squared", C.c_float), ("velocity_multiply", C.c_int32)]


class _Results(C.Structure):
    _fields_ = [("cost", C.POINTER(C.c_uint64)), ("flags", C.POINTER(C.c_uint32)), ("potted", C.POINTER(C.c_uint32)),
                ("final_time", C.POINTER(C.c_float)), ("final_pos", C.POINTER(C.c_float)),
                ("event_log", C.POINTER(C.c_int32))]


_lib = None


def load_library(path=None):
    """Loads libbilliard_cuda.so; raises BcuError when it has not been built (no fallback)."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise BcuError(f"{p} is missing: build it with `python -m billiard_ai_b200.build` "
                       "(this package has no CPU path)")
    lib = C.CDLL(p)
    lib.bcu_version.restype = C.c_char_p
    lib.bcu_last_error.restype = C.c_char_p
    lib.bcu_last_error.argtypes = [C.c_void_p]
    lib.bcu_launch_count.restype = C.c_int64
    lib.bcu_launch_count.argtypes = [C.c_void_p]
    lib.bcu_destroy.argtypes = [C.c_void_p]
    lib.bcu_destroy.restype = None
    lib.bcu_host_alloc.argtypes = [C.c_size_t, C.POINTER(C.c_void_p)]
    lib.bcu_host_free.argtypes = [C.c_void_p]
    lib.bcu_get_prune.argtypes = [C.c_void_p]
    lib.bcu_test_merge_records.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]
    lib.bcu_set_planner.argtypes = [C.c_void_p, C.c_void_p]
    lib.bcu_plan.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
    lib.bcu_fetch_plan.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.POINTER(C.c_int32), C.POINTER(C.c_float), C.POINTER(C.c_float),
                                   C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.bcu_evaluate_plan.argtypes = [C.c_void_p, C.c_void_p]
    lib.bcu_evaluate_plan_range.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_void_p]
    lib.bcu_engine_rank_info.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.bcu_search_levels.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_float), C.POINTER(C.c_int32), C.c_void_p, C.c_void_p,
                                      C.c_void_p]
    if path is None:
        _lib = lib
    return lib


def topk_static(d_cost, d_flags, n, first_candidate, k, d_out_cost, d_out_id, d_out_count, stream=0, lib=None):
    """bcu_topk_static (include/billiard_cuda.h): top-k of (cost, id) among VALID candidates on DEVICE pointers (integers),
    queued on `stream`, no allocation anywhere (static device scratch, hence n <= 2^22 and k <= 1024).  Engine-free; does not
    synchronise.  Raises BcuError with the library's message."""
    lib = lib or load_library()
    err = C.create_string_buffer(256)
    rc = lib.bcu_topk_static(C.c_void_p(d_cost), C.c_void_p(d_flags), C.c_int64(n), C.c_int64(first_candidate), C.c_int32(k),
                             C.c_void_p(d_out_cost), C.c_void_p(d_out_id), C.c_void_p(d_out_count), C.c_void_p(stream), err,
                             C.c_int32(len(err)))
    if rc != 0:
        raise BcuError(f"bcu_topk_static failed ({rc}): {err.value.decode()}")


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _null(t):
    return C.POINTER(t)()


class _LazyResults(dict):
    """Per-strike results; "flags" (status bits) and "n_events" are unpacked from the raw flag words on first use."""

    def __missing__(self, key):
        if key == "flags":
            v = self["flags_raw"] & 0xff
        elif key == "n_events":
            v = ((self["flags_raw"] >> 8) & 0xff).astype(np.int32)
        else:
            raise KeyError(key)
        self[key] = v
        return v

    def _all(self):
        self["flags"]; self["n_events"]
        return self

    def keys(self):
        return dict.keys(self._all())

    def items(self):
        return dict.items(self._all())

    def values(self):
        return dict.values(self._all())

    def __iter__(self):
        return dict.__iter__(self._all())

    def __contains__(self, key):
        return key in ("flags", "n_events") or dict.__contains__(self, key)

    def get(self, key, default=None):
        return self[key] if key in self else default


class Engine:
    """One engine = one CUDA device + one stream + the table / rules constants."""

    def __init__(self, rails6, pockets3, ball_radius, scores, cue_type=0, device=0, lib_path=None):
        self.lib = load_library(lib_path)   # lib_path: a tuning / guard-variant build (billiard_ai_b200/build.py)
        self._rails = np.ascontiguousarray(rails6, dtype=np.float32)
        self._pockets = np.ascontiguousarray(np.asarray(pockets3, dtype=np.float32)[:, :3])
        self._scores = np.ascontiguousarray(scores, dtype=np.float64)
        td = _TableDesc(len(self._rails), len(self._pockets), C.c_float(np.float32(ball_radius)),
                        _p(self._rails, C.c_float), _p(self._pockets, C.c_float), _null(C.c_float))
        rd = _RulesDesc(len(self._scores), int(cue_type), _p(self._scores, C.c_double))
        h = C.c_void_p()
        rc = self.lib.bcu_create(C.byref(td), C.byref(rd), C.c_int(device), C.byref(h))
        if rc != 0:
            raise BcuError(f"bcu_create failed: {_STATUS.get(rc, rc)} (no CPU fallback)")
        self.h = h
        self.n_rails = len(self._rails)
        self.n_balls = 0
        self.n_states = 0
        self._last_n = 0
        self._beam_k = 0
        self._pinned = []
        self._result_cache = {}

    @classmethod
    def from_table(cls, tbl, scores, cue_type=0, device=0, lib_path=None):
        """tbl: dict in the layout of tests/golden/table.json (reference configuration.json)."""
        rails6 = np.array([[r["start"][0], r["start"][1], r["end"][0], r["end"][1], r["shift"][0], r["shift"][1]]
                           for r in tbl["rails"]], dtype=np.float32)
        pockets3 = np.array([[p["position"][0], p["position"][1], p["radius"]] for p in tbl["pockets"]],
                            dtype=np.float32)
        return cls(rails6, pockets3, tbl["ballRadius"], scores, cue_type, device, lib_path=lib_path)

    def close(self):
        if getattr(self, "h", None):
            self.lib.bcu_destroy(self.h)
            self.h = None
            self._result_cache = {}
            for ptr in self._pinned:
                self.lib.bcu_host_free(ptr)
            self._pinned = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc, what):
        if rc != 0:
            msg = self.lib.bcu_last_error(self.h)
            raise BcuError(f"{what}: {_STATUS.get(rc, rc)}: {msg.decode() if msg else ''}")

    # ---- table ----------------------------------------------------------------------------------
    def rail_derived(self):
        out = np.zeros((self.n_rails, 6), dtype=np.float32)
        self._ck(self.lib.bcu_get_rail_derived(self.h, _p(out, C.c_float)), "bcu_get_rail_derived")
        return out

    # ---- states ---------------------------------------------------------------------------------
    def set_states(self, pos, types, expected_mask, base_cost=None):
        pos = np.ascontiguousarray(pos, dtype=np.float32)
        types = np.ascontiguousarray(types, dtype=np.int32)
        if types.ndim == 1:
            types = types[None]
            pos = pos[None]
        ns, nb = types.shape
        em = np.ascontiguousarray(np.broadcast_to(np.asarray(expected_mask, dtype=np.uint32), (ns,)))
        bc = None if base_cost is None else np.ascontiguousarray(
            np.broadcast_to(np.asarray(base_cost, dtype=np.uint64), (ns,)))
        self._ck(self.lib.bcu_set_states(self.h, C.c_int32(ns), C.c_int32(nb), _p(pos, C.c_float), _p(types, C.c_int32),
                                         _p(em, C.c_uint32), _p(bc, C.c_uint64) if bc is not None else _null(C.c_uint64)),
                 "bcu_set_states")
        self.n_states, self.n_balls = ns, nb

    # ---- explicit candidates ----------------------------------------------------------------------
    def pinned_array(self, shape, dtype):
        """numpy array over page-locked host memory (bcu_host_alloc); freed when the engine is closed."""
        dtype = np.dtype(dtype)
        count = int(np.prod(shape))
        ptr = C.c_void_p()
        self._ck(self.lib.bcu_host_alloc(C.c_size_t(max(1, count) * dtype.itemsize), C.byref(ptr)), "bcu_host_alloc")
        self._pinned.append(ptr)
        buf = (C.c_char * (max(1, count) * dtype.itemsize)).from_address(ptr.value)
        return np.frombuffer(buf, dtype=dtype, count=count).reshape(shape)

    def _result_buffers(self, n, final_pos, event_log, reuse=False):
        if reuse:   # one cached set of pinned buffers per shape: the previous call's results are overwritten
            key = (n, self.n_balls, bool(final_pos), bool(event_log))
            if key not in self._result_cache:
                self._result_cache[key] = dict(
                    cost=self.pinned_array(n, np.uint64), flags=self.pinned_array(n, np.uint32),
                    potted=self.pinned_array(n, np.uint32), final_time=self.pinned_array(n, np.float32),
                    final_pos=self.pinned_array((n, self.n_balls, 2), np.float32) if final_pos else None,
                    event_log=self.pinned_array((n, 32, 4), np.int32) if event_log else None)
            out = dict(self._result_cache[key])
        else:
            out = dict(cost=np.zeros(n, np.uint64), flags=np.zeros(n, np.uint32), potted=np.zeros(n, np.uint32),
                       final_time=np.zeros(n, np.float32),
                       final_pos=np.zeros((n, self.n_balls, 2), np.float32) if final_pos else None,
                       event_log=np.full((n, 32, 4), -1, np.int32) if event_log else None)
        res = _Results(_p(out["cost"], C.c_uint64), _p(out["flags"], C.c_uint32), _p(out["potted"], C.c_uint32),
                       _p(out["final_time"], C.c_float),
                       _p(out["final_pos"], C.c_float) if final_pos else _null(C.c_float),
                       _p(out["event_log"], C.c_int32) if event_log else _null(C.c_int32))
        return out, res

    @staticmethod
    def _finish(out):
        res = _LazyResults(out)
        res["flags_raw"] = res.pop("flags")   # status bits 0-7, number of events bits 8-15 (include/billiard_cuda.h)
        return res

    def evaluate(self, cand_state, vx, vy, search_cost, base_cost, final_pos=True, event_log=False, reuse_buffers=False,
                 select_k=None):
        """reuse_buffers=True returns views of cached page-locked result buffers (overwritten by the next such call).
        select_k: also select the k best valid strikes (bcu_evaluate_level); fetch them with fetch_beam()."""
        vx = np.ascontiguousarray(vx, dtype=np.float32)
        vy = np.ascontiguousarray(vy, dtype=np.float32)
        n = len(vx)
        cs = np.ascontiguousarray(np.broadcast_to(np.asarray(cand_state, dtype=np.int32), (n,)))
        sc = np.ascontiguousarray(np.broadcast_to(np.asarray(search_cost, dtype=np.uint64), (n,)))
        bc = np.ascontiguousarray(np.broadcast_to(np.asarray(base_cost, dtype=np.uint64), (n,)))
        out, res = self._result_buffers(n, final_pos, event_log, reuse=reuse_buffers)
        if select_k:
            self._ck(self.lib.bcu_evaluate_level(self.h, C.c_int64(n), _p(cs, C.c_int32), _p(vx, C.c_float), _p(vy, C.c_float),
                                                 _p(sc, C.c_uint64), _p(bc, C.c_uint64), C.byref(res), C.c_int32(select_k)),
                     "bcu_evaluate_level")
            self._beam_k = select_k
        else:
            self._ck(self.lib.bcu_evaluate(self.h, C.c_int64(n), _p(cs, C.c_int32), _p(vx, C.c_float), _p(vy, C.c_float),
                                           _p(sc, C.c_uint64), _p(bc, C.c_uint64), C.byref(res)), "bcu_evaluate")
        self._last_n = n
        return self._finish(out)

    # ---- direction x speed grid -------------------------------------------------------------------
    def evaluate_grid(self, dirs, speeds, search_cost=100000, first=0, count=-1):
        dirs = np.ascontiguousarray(dirs, dtype=np.float32)
        speeds = np.ascontiguousarray(speeds, dtype=np.float32)
        total = self.n_states * len(dirs) * len(speeds)
        self._ck(self.lib.bcu_evaluate_grid(self.h, C.c_int32(len(dirs)), C.c_int32(len(speeds)), _p(dirs, C.c_float),
                                            _p(speeds, C.c_float), C.c_uint64(search_cost), C.c_int64(first),
                                            C.c_int64(count)), "bcu_evaluate_grid")
        self._last_n = total - first if count < 0 else count
        return self._last_n

    def rerun(self):
        self._ck(self.lib.bcu_rerun(self.h), "bcu_rerun")

    def fetch(self, offset=0, n=None, final_pos=True):
        n = self._last_n - offset if n is None else n
        out, res = self._result_buffers(n, final_pos, False)
        self._ck(self.lib.bcu_fetch_results(self.h, C.c_int64(offset), C.c_int64(n), C.byref(res)), "bcu_fetch_results")
        return self._finish(out)

    def topk(self, k):
        cost = np.zeros(k, np.uint64)
        ids = np.zeros(k, np.uint64)
        cnt = C.c_int32(0)
        self._ck(self.lib.bcu_topk(self.h, C.c_int32(k), _p(cost, C.c_uint64), _p(ids, C.c_uint64), C.byref(cnt)),
                 "bcu_topk")
        return cost[:cnt.value], ids[:cnt.value]

    def gather(self, ids):
        ids = np.ascontiguousarray(ids, dtype=np.uint64)
        n = len(ids)
        pos = np.zeros((n, self.n_balls, 2), np.float32)
        potted = np.zeros(n, np.uint32)
        tm = np.zeros(n, np.float32)
        self._ck(self.lib.bcu_gather(self.h, C.c_int32(n), _p(ids, C.c_uint64), _p(pos, C.c_float),
                                     _p(potted, C.c_uint32), _p(tm, C.c_float)), "bcu_gather")
        return pos, potted, tm

    # ---- one break level, device resident ------------------------------------------------------------
    def level_select(self, k):
        self._ck(self.lib.bcu_level_select(self.h, C.c_int32(k)), "bcu_level_select")
        self._beam_k = k

    def level_rerun(self, k):
        self._ck(self.lib.bcu_level_rerun(self.h, C.c_int32(k)), "bcu_level_rerun")
        self._beam_k = k

    def fetch_beam(self, with_pos=True):
        k = self._beam_k
        cost = np.zeros(k, np.uint64)
        ids = np.zeros(k, np.uint64)
        potted = np.zeros(k, np.uint32)
        tm = np.zeros(k, np.float32)
        pos = np.zeros((k, self.n_balls, 2), np.float32) if with_pos else None
        cnt = C.c_int32(0)
        self._ck(self.lib.bcu_fetch_beam(self.h, C.byref(cnt), _p(cost, C.c_uint64), _p(ids, C.c_uint64),
                                         _p(potted, C.c_uint32), _p(tm, C.c_float),
                                         _p(pos, C.c_float) if with_pos else _null(C.c_float)), "bcu_fetch_beam")
        n = cnt.value
        return dict(cost=cost[:n], id=ids[:n], potted=potted[:n], final_time=tm[:n],
                    final_pos=pos[:n] if with_pos else None)

    def set_id_base(self, base):
        self._ck(self.lib.bcu_set_id_base(self.h, C.c_int64(base)), "bcu_set_id_base")

    def sync(self):
        self._ck(self.lib.bcu_sync(self.h), "bcu_sync")

    # ---- multi-GPU -----------------------------------------------------------------------------------
    @staticmethod
    def nccl_library_path():
        """torch's bundled NCCL (the same library torch.distributed uses in this process)."""
        try:
            import nvidia.nccl
            d = os.path.join(os.path.dirname(nvidia.nccl.__path__[0] + "/"), "lib")
            p = os.path.join(d, "libnccl.so.2")
            if os.path.exists(p):
                return p
        except Exception:
            pass
        return "libnccl.so.2"

    def comm_unique_id(self):
        buf = (C.c_char * 128)()
        rc = self.lib.bcu_comm_unique_id(self.nccl_library_path().encode(), buf)
        if rc != 0:
            raise BcuError("bcu_comm_unique_id failed (cannot load NCCL)")
        return bytes(buf)

    def comm_init(self, n_ranks, rank, unique_id=None):
        uid = (C.c_char * 128).from_buffer_copy(unique_id) if unique_id is not None else None
        self._ck(self.lib.bcu_comm_init(self.h, C.c_int32(n_ranks), C.c_int32(rank), uid,
                                        self.nccl_library_path().encode()), "bcu_comm_init")

    # ---- single strike with layer trace -------------------------------------------------------------
    def simulate_trace(self, pos, types, vx, vy):
        pos = np.ascontiguousarray(pos, dtype=np.float32)
        types = np.ascontiguousarray(types, dtype=np.int32)
        n = len(types)
        times = np.zeros(MAX_LAYERS, np.float32)
        rec = np.zeros((MAX_LAYERS, n, TRACE_REC), np.float32)
        nl = C.c_int32(0)
        self._ck(self.lib.bcu_simulate_trace(self.h, C.c_int32(n), _p(pos, C.c_float), _p(types, C.c_int32),
                                             C.c_float(vx), C.c_float(vy), C.byref(nl), _p(times, C.c_float),
                                             _p(rec, C.c_float)), "bcu_simulate_trace")
        return nl.value, times[:nl.value], rec[:nl.value]

    def trace_strikes(self, cand_state, vx, vy):
        """Full per-layer traces of n strikes on the current states -> (layers [n], times [n][32], rec [n][32][balls][10]);
        layers[i] == 0 where the reference's simulate() would return nullopt."""
        vx = np.ascontiguousarray(vx, dtype=np.float32)
        vy = np.ascontiguousarray(vy, dtype=np.float32)
        n = len(vx)
        cs = np.ascontiguousarray(np.broadcast_to(np.asarray(cand_state, dtype=np.int32), (n,)))
        layers = np.zeros(n, np.int32)
        times = np.zeros((n, MAX_LAYERS), np.float32)
        rec = np.zeros((n, MAX_LAYERS, self.n_balls, TRACE_REC), np.float32)
        self._ck(self.lib.bcu_trace_strikes(self.h, C.c_int32(n), _p(cs, C.c_int32), _p(vx, C.c_float), _p(vy, C.c_float),
                                            _p(layers, C.c_int32), _p(times, C.c_float), _p(rec, C.c_float)),
                 "bcu_trace_strikes")
        return layers, times, rec

    # ---- measurement hooks ------------------------------------------------------------------------
    def last_kernel_ms(self):
        ms = C.c_float(0)
        self._ck(self.lib.bcu_last_kernel_ms(self.h, C.byref(ms)), "bcu_last_kernel_ms")
        return ms.value

    def timer_begin(self):
        self._ck(self.lib.bcu_timer_begin(self.h), "bcu_timer_begin")

    def timer_end(self):
        ms = C.c_float(0)
        self._ck(self.lib.bcu_timer_end(self.h, C.byref(ms)), "bcu_timer_end")
        return ms.value

    def flush_l2(self):
        self._ck(self.lib.bcu_flush_l2(self.h), "bcu_flush_l2")

    def issue_rate(self, kind):
        out = C.c_double(0)
        self._ck(self.lib.bcu_issue_rate(self.h, C.c_int32(kind), C.byref(out)), "bcu_issue_rate")
        return out.value

    def enable_counters(self, on=True):
        self._ck(self.lib.bcu_enable_counters(self.h, C.c_int(int(on))), "bcu_enable_counters")

    def counters(self):
        out = np.zeros(5, np.uint64)
        self._ck(self.lib.bcu_get_counters(self.h, _p(out, C.c_uint64)), "bcu_get_counters")
        return dict(zip(["events", "pair_items", "quartics", "rail_items", "pocket_items"], out.tolist()))

    def set_prune(self, mode):
        """0 off (default: every quartic the reference solves is solved), 1 on (opt-in: audited, not proven), 2 audit;
        see include/billiard_cuda.h."""
        self._ck(self.lib.bcu_set_prune(self.h, C.c_int(int(mode))), "bcu_set_prune")

    def get_prune(self):
        return int(self.lib.bcu_get_prune(self.h))

    def test_merge_records(self, blocks, k, n_balls):
        """Test hook: G host record blocks (uint8 arrays, layout of include/billiard_cuda.h: bcu_test_merge_records) -> the merged block, computed
        by the device merge kernel that bcu_level_select runs after the all-gather."""
        blocks = [np.ascontiguousarray(b, dtype=np.uint8) for b in blocks]
        cat = np.ascontiguousarray(np.concatenate(blocks))
        out = np.zeros(len(blocks[0]), np.uint8)
        self._ck(self.lib.bcu_test_merge_records(self.h, C.c_int32(len(blocks)), C.c_int32(k), C.c_int32(n_balls),
                                                 cat.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)),
                 "bcu_test_merge_records")
        return out

    def prune_counters(self):
        out = np.zeros(2, np.uint64)
        self._ck(self.lib.bcu_get_prune_counters(self.h, _p(out, C.c_uint64)), "bcu_get_prune_counters")
        return dict(zip(["skipped", "audit_failures"], out.tolist()))

    def launch_count(self):
        return int(self.lib.bcu_launch_count(self.h))

    # ---- the reference's geometric planner on the device (include/billiard_cuda.h: bcu_set_planner ... bcu_evaluate_plan) ----
    def set_planner(self, tbl, velocity_multiply=True):
        """tbl: dict in the layout of tests/golden/table.json, the table this engine was created with (same rail order);
        needs its `railMapOrder` / `railSetOrder` (the reference's container iteration orders) and the pockets' potting points."""
        ids = [r["id"] for r in tbl["rails"]]
        self._plan_start = np.ascontiguousarray([[r["start"][0], r["start"][1]] for r in tbl["rails"]], np.float32)
        self._plan_order = np.ascontiguousarray([ids.index(i) for i in tbl["railMapOrder"]], np.int32)
        self._plan_set = np.ascontiguousarray([ids.index(i) for i in tbl["railSetOrder"]], np.int32)
        self._plan_pockets = np.ascontiguousarray([[p["pottingPoint"][0], p["pottingPoint"][1], p["normal"][0], p["normal"][1]]
                                                   for p in tbl["pockets"]], np.float32)
        length, width = np.float32(tbl["innerTableLength"]), np.float32(tbl["innerTableWidth"])
        diag = np.float32(np.float32(length * length) + np.float32(width * width))   # util.cpp:73, binary32
        d = _PlannerDesc(len(ids), _p(self._plan_start, C.c_float), _p(self._plan_order, C.c_int32), len(self._plan_set),
                         _p(self._plan_set, C.c_int32), len(self._plan_pockets), _p(self._plan_pockets, C.c_float),
                         C.c_float(np.float32(tbl["minimalPocketVelocity"])), C.c_float(diag), int(bool(velocity_multiply)))
        self._ck(self.lib.bcu_set_planner(self.h, C.byref(d)), "bcu_set_planner")

    def plan(self):
        """Plans every state of the last set_states on the device -> number of strikes planned."""
        n = C.c_int64(0)
        self._ck(self.lib.bcu_plan(self.h, C.byref(n)), "bcu_plan")
        self._plan_n = n.value
        return n.value

    def fetch_plan(self):
        n = self._plan_n
        out = dict(state=np.zeros(n, np.int32), vx=np.zeros(n, np.float32), vy=np.zeros(n, np.float32),
                   search_cost=np.zeros(n, np.uint64), cost=np.zeros(n, np.uint64))
        if n:
            self._ck(self.lib.bcu_fetch_plan(self.h, 0, n, _p(out["state"], C.c_int32), _p(out["vx"], C.c_float), _p(out["vy"], C.c_float),
                                             _p(out["search_cost"], C.c_uint64), _p(out["cost"], C.c_uint64)), "bcu_fetch_plan")
        return out

    def evaluate_plan(self, final_pos=True, event_log=False):
        """Simulates the planned strikes (as evaluate() on them)."""
        n = self._plan_n
        out, res = self._result_buffers(n, final_pos, event_log)
        if n:
            self._ck(self.lib.bcu_evaluate_plan(self.h, C.byref(res)), "bcu_evaluate_plan")
        self._last_n = n
        return self._finish(out)

    def rank_info(self):
        """(rank, number of ranks) of the engine's communicator."""
        r, n = C.c_int32(0), C.c_int32(1)
        self._ck(self.lib.bcu_engine_rank_info(self.h, C.byref(r), C.byref(n)), "bcu_engine_rank_info")
        return r.value, n.value
