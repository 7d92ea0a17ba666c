#!/usr/bin/env python
"""bench.py -- strikes/s of the strike-evaluation hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" is one break level of the search over one batch of synthetic candidates (BASELINE.json configs[1],
restated in SURVEY.md 8d "Config 2"): the 22-ball snooker layout x (1024 directions x 64 speeds) = 65 536
candidate strikes PER GPU (weak scaling: N GPUs evaluate N x 1024 directions, sharded by contiguous id range);
each strike is simulated to rest, costed, the k = 1024 best valid strikes are selected on the device and -- for
N > 1 -- merged across GPUs with one ncclAllGather of the per-GPU record blocks.

  value : whole-job strikes/s, inputs resident in HBM (CUDA events on the engine's stream, max over ranks), with the
          reach bound ON (bcu_set_prune 1: audited, not proven -- DESIGN.md section 4)
  value_exact : the same with the reach bound OFF (the C ABI's default: every test and every quartic of the reference)
  e2e / e2e_exact : the same level through the host-buffer C-ABI calls (bcu_set_states + bcu_evaluate_level (= bcu_evaluate
          + bcu_level_select) + bcu_fetch_beam): pinned host candidate arrays in; per-candidate cost / flags / potted mask /
          final time and the beam (k best strikes with their final table positions) out, wall clock
  parity : every strike of the timed workload (all N x 65 536) against the reference's own code on the host cores, once per
          run, untimed: mismatches, best strike, and -- N > 1 -- the merged beam against one GPU evaluating the whole grid
  roofline     : the pipeline's FP64 work / its CUDA-event time vs. the FP64 vector issue peak measured in the same run
                 (un-fused DADD+DMUL: FMA contraction is forbidden on this path); three fractions: algorithmic (the
                 reference's quartic count), executed (quartics actually solved), issue-slot utilisation
  cpu_baseline : the reference's own CPU code (oracle/_ref/libref_oracle.so) on the whole grid of one GPU, all host threads
  stats, extra : a longer timed region (>= 200 steps: mean / median / p95) and secondary workloads (16-ball level, config 5
                 slice, multi-break searches), each with both prune modes

--impl reference times the reference CPU implementation only (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_DIRS, N_SPEEDS, BEAM = 1024, 64, 1024
METRIC = "simulated_strikes_per_sec"
UNIT = "strikes/s"
# SURVEY.md 8(d) working per-call costs (FLOP): pair pre-check, rail test, pocket test, per-ball advance; quartic FP64
W_PAIR, W_RAIL, W_POCKET, W_ADVANCE, W_QUARTIC_FP64 = 25.0, 20.0, 20.0, 12.0, 1000.0


def workload_desc(world):
    return {"workload": "config2: snooker 22-ball layout (open table, seed 22), 1-break search, 65536 strikes per GPU "
                        f"({N_DIRS}x{world} directions x {N_SPEEDS} speeds), beam {BEAM}",
            "n_balls": 22, "candidates_per_gpu": N_DIRS * N_SPEEDS, "candidates_total": N_DIRS * N_SPEEDS * world,
            "beam": BEAM, "l2": "flushed between timed steps (256 MB write, untimed)",
            "parallelism": f"candidate shards x{world} (contiguous id ranges, directions enumerated round-robin over the shards), one ncclAllGather of top-k records per level" if world > 1
                           else "single GPU"}


# ---- clocks ---------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.idx), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc:
            time.sleep(0.12)
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm)}


# ---- reference CPU arm ----------------------------------------------------------------------------------------
def load_reference():
    """oracle/_ref/libref_oracle.so = the reference's own sources (kind "reference"); the only place outside
    tests/ and smoke() that executes anything under oracle/."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    if os.path.exists(oracle_lib.REF_SO):
        return oracle_lib.RefOracle(), "reference"
    return oracle_lib.PortOracle(), "port"


def sharded_grid(world):
    """The N-GPU workload: 1024 x N directions x 64 speeds.  Candidate ids are sharded by contiguous range, so
    the direction table is enumerated shard by shard -- shard r holds the directions theta_k with k = r (mod N) -- and every
    GPU sweeps the whole circle instead of one sector (a 90-degree sector of this layout is 15 % more work than the
    average, and a level takes as long as its slowest shard).  One GPU: the plain table."""
    from billiard_ai_b200 import workloads as W
    dirs, speeds = W.grid(N_DIRS * world, N_SPEEDS)
    order = np.arange(N_DIRS * world).reshape(N_DIRS, world).T.reshape(-1)
    return np.ascontiguousarray(dirs[order]), speeds


def reference_pass(oracle, world, threads, shard=None):
    """One pass of the reference's CPU code over the candidates of the bench workload -- the same grid, the same ids --
    either all N x 65 536 of them (shard None: the parity check) or one GPU's shard of 65 536 (the timed CPU baseline).
    -> (results, seconds)"""
    from billiard_ai_b200 import workloads as W
    pos, types = W.snooker22_open()
    dirs, speeds = sharded_grid(world)
    vx, vy = W.grid_velocities(dirs, speeds)
    if shard is not None:
        lo, hi = W.shard_range(len(vx), shard, world)
        vx, vy = vx[lo:hi], vy[lo:hi]
    t0 = time.perf_counter()
    out = oracle.evaluate(pos[None], types[None], 1 << W.TYPE_CODE["RED"], 0, vx, vy, W.GRID_SEARCH_COST, 0, threads=threads)
    return out, time.perf_counter() - t0


def run_reference(args, rank, world):
    """The reference arm: the reference's own CPU implementation of the path on the box's host cores.  A step is one pass
    over ONE GPU's shard of the workload (all 65 536 candidates of shard 0: at N = 1 that is the whole workload)."""
    if rank != 0:
        return
    oracle, kind = load_reference()
    threads = os.cpu_count() or 1
    for _ in range(min(args.warmup, 2)):
        reference_pass(oracle, world, threads, shard=0)
    tot_t, tot_n, tot_ev = 0.0, 0, 0
    for _ in range(args.steps):
        out, dt = reference_pass(oracle, world, threads, shard=0)
        tot_t += dt; tot_n += len(out["cost"]); tot_ev += int(out["n_events"].sum())
    value = tot_n / tot_t
    sample = (f"every step = all {N_DIRS * N_SPEEDS} candidates of shard 0 of {world} (the same grid and ids the GPU arm "
              f"evaluates; {N_DIRS * N_SPEEDS * world} in the whole job)")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32+f64", "data": "synthetic",
            "config": workload_desc(world), "events_per_sec": tot_ev / tot_t,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ---- B200 arm -------------------------------------------------------------------------------------------------
def measured_op_counts():
    """Per-kernel FP instruction counts of one level of this workload from the committed ncu pass
    (profiles/r02_level_fp_ops.json, tools/ncu_fp_ops.py): replaces SURVEY.md's working figure of 1000 FP64 FLOP per
    quartic by what the quartic kernel executes.  None when the file is absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_level_fp_ops.json")) as f:
            return json.load(f)
    except Exception:
        return None


def compare_with_reference(got, want, tie):
    """-> (tie-free strikes compared, mismatching strikes): integers equal, floats bit-equal."""
    n = len(want["cost"])
    clean = tie["min_gap"] >= 1e-9 if tie is not None and tie.get("min_gap") is not None else np.ones(n, bool)
    bad = np.zeros(n, bool)
    for key in ("cost", "flags", "n_events", "potted"):
        bad |= got[key] != want[key]
    bad |= got["final_time"].view(np.uint32) != want["final_time"].view(np.uint32)
    bad |= (got["final_pos"].view(np.uint32) != want["final_pos"].view(np.uint32)).reshape(n, -1).any(axis=1)
    return int(clean.sum()), int((bad & clean).sum()), clean


def timed_levels(eng, steps, barrier):
    """`steps` levels, L2 flushed before each, CUDA events on the engine's stream -> per-step ms, per-step kernel ms."""
    step_ms, kernel_ms = [], []
    barrier()
    for _ in range(steps):
        eng.flush_l2()
        eng.timer_begin()
        eng.level_rerun(BEAM)
        step_ms.append(eng.timer_end())
        kernel_ms.append(eng.last_kernel_ms())
    barrier()
    return step_ms, kernel_ms


def run_b200(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from billiard_ai_b200 import Engine, workloads as W

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    eng = Engine.from_table(W.default_table(), W.SNOOKER_SCORES, cue_type=W.TYPE_CODE["WHITE"], device=local_rank)
    if world > 1:
        box = [eng.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        eng.comm_init(world, rank, box[0])

    pos, types = W.snooker22_open()
    red = 1 << W.TYPE_CODE["RED"]
    dirs, speeds = sharded_grid(world)
    total = len(dirs) * len(speeds)
    lo, hi = W.shard_range(total, rank, world)
    mine = hi - lo

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        eng.sync()

    def allmax(x):
        if world == 1:
            return float(x)
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def resident_launch(prune):
        eng.set_prune(prune)
        eng.set_states(pos, types, red)
        eng.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST, first=lo, count=mine)   # uploads + first launch
        for _ in range(max(args.warmup, 3)):
            eng.level_rerun(BEAM)
        eng.sync()

    # ---- timed: value (resident inputs, reach bound on) ------------------------------------------------------------
    resident_launch(1)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    # a few more untimed levels right in front of the timed region, on every rank, after the clock sampler has started: a
    # 4-GPU run showed a first region at 2.9 ms per step (one rank slow for ~20 steps after start-up) in front of a 200-step
    # region at 1.65 -- transients of process start-up belong to the warm-up, not to the K timed steps
    SETTLE = 16
    for _ in range(SETTLE):
        eng.flush_l2()
        eng.level_rerun(BEAM)
    barrier()
    launches0 = eng.launch_count()
    wall0 = time.perf_counter()
    step_ms, kernel_ms = timed_levels(eng, args.steps, barrier)
    wall = time.perf_counter() - wall0
    launches = eng.launch_count() - launches0
    t_ms = allmax(float(np.sum(step_ms)))
    # a longer region on the same launch: spread of the step time (the K-step region above is the contract's number)
    long_steps = max(200, args.steps)
    long_ms, _ = timed_levels(eng, long_steps, barrier)
    clocks = sampler.stop() if rank == 0 else None
    long_sum = allmax(float(np.sum(long_ms)))
    beam = eng.fetch_beam()

    # ---- work counters (untimed extra launches) -> events/s and the roofline's work ---------------------------------
    eng.enable_counters(True)
    eng.rerun()
    cnt = eng.counters()
    eng.set_prune(2)            # audit launch: every quartic the reference solves is solved and counted
    eng.rerun()
    cnt_ref = eng.counters()
    audit = eng.prune_counters()
    assert audit["audit_failures"] == 0, audit
    eng.set_prune(1)
    eng.enable_counters(False)
    cvec = torch.tensor([cnt["events"], cnt["pair_items"], cnt_ref["quartics"], cnt["rail_items"], cnt["pocket_items"],
                         cnt["quartics"]], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(cvec)
    ev_total = float(cvec[0].item())

    # ---- timed: value_exact (reach bound off: the C ABI's default) ----------------------------------------------------
    resident_launch(0)
    exact_ms, exact_kernel_ms = timed_levels(eng, args.steps, barrier)
    t_exact_ms = allmax(float(np.sum(exact_ms)))
    beam_exact = eng.fetch_beam()
    assert np.array_equal(beam_exact["cost"], beam["cost"]) and np.array_equal(beam_exact["id"], beam["id"]), \
        "reach bound on / off select different beams"

    # ---- e2e: host buffers through the C ABI ---------------------------------------------------------------------------
    vx_all, vy_all = W.grid_velocities(dirs, speeds)

    def pinned(a):
        return torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    keep = [pinned(vx_all[lo:hi]), pinned(vy_all[lo:hi]), pinned(np.zeros(mine, np.int32)),
            pinned(np.full(mine, W.GRID_SEARCH_COST, np.int64)), pinned(np.zeros(mine, np.int64))]
    hvx, hvy, hst, hsc, hbc = [t.numpy() for t in keep]
    hsc = hsc.view(np.uint64); hbc = hbc.view(np.uint64)
    eng.set_id_base(lo)

    def e2e_step():
        eng.set_states(pos, types, red)
        res = eng.evaluate(hst, hvx, hvy, hsc, hbc, final_pos=False, event_log=False, reuse_buffers=True, select_k=BEAM)
        b = eng.fetch_beam()
        return res, b
    e2e_steps = max(1, min(args.steps, 10))

    def e2e_run(prune):
        eng.set_prune(prune)
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            res, b2 = e2e_step()
        barrier()
        secs = allmax(time.perf_counter() - t0)
        assert np.array_equal(b2["cost"], beam["cost"]) and np.array_equal(b2["id"], beam["id"]), "e2e beam != resident beam"
        return secs
    e2e_s = e2e_run(1)
    e2e_exact_s = e2e_run(0)
    nb = len(types)
    h2d = mine * (4 + 4 + 4 + 8 + 8) + nb * (8 + 4) + 4
    d2h = mine * (8 + 4 + 4 + 4) + BEAM * (24 + nb * 8)   # per-strike cost / flags / potted / time + the beam records

    # ---- config 4 (and config 3 on one GPU) through bcu_search_levels: every rank takes part (one allgather per level) ------
    searches = None
    if not args.no_extra:
        searches = search_workloads(eng, W, rank, world, dist if world > 1 else None)
        eng.set_prune(1)

    if rank != 0:
        barrier()           # rank 0 checks parity on its own GPU meanwhile (no collective in there)
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- parity: every strike of the timed workload against the reference's own code, untimed ----------------------------
    #      (rank 0 evaluates the WHOLE N x 65 536 grid on its own GPU, so at N > 1 the merged beam of the N ranks is checked
    #      against a single-GPU evaluation as well as against the reference)
    parity, cpu = None, None
    if not args.no_cpu_baseline:
        oracle, kind = load_reference()
        import oracle_lib
        threads = os.cpu_count() or 1
        want, _ = reference_pass(oracle, world, threads)
        tie = None
        if os.path.exists(oracle_lib.PORT_SO):   # near-tie information (the port is bit-identical to the reference build)
            tie, _ = reference_pass(oracle_lib.PortOracle(), world, threads)
        parity = {"reference": kind, "compared": 0, "mismatches": 0}
        for prune in (1, 0):
            eng.set_prune(prune)
            eng.set_states(pos, types, red)
            eng.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST, first=0, count=total)
            got = eng.fetch()
            compared, bad, clean = compare_with_reference(got, want, tie)
            parity["compared"] += compared if prune == 1 else 0
            parity["mismatches"] += bad
            parity[f"mismatches_prune{prune}"] = bad
        ids = np.nonzero((want["flags"] & 2) != 0)[0]
        order = ids[np.lexsort((ids, want["cost"][ids]))][:BEAM]
        tc, ti = eng.topk(BEAM)     # the whole grid on this one GPU, no collective
        parity["near_ties"] = int((~clean).sum())
        parity["best_strike_equal"] = bool(len(order) and len(beam["id"]) and int(beam["id"][0]) == int(order[0])
                                           and int(beam["cost"][0]) == int(want["cost"][order[0]]))
        parity["beam_equals_reference_topk"] = bool(np.array_equal(beam["id"], order.astype(np.uint64))
                                                    and np.array_equal(beam["cost"], want["cost"][order]))
        parity["merged_beam_equals_single_gpu"] = bool(np.array_equal(beam["id"], ti) and np.array_equal(beam["cost"], tc))
        gpos, gpot, _ = eng.gather(beam["id"])
        parity["beam_end_states_equal"] = bool(np.array_equal(gpos.view(np.uint32), beam["final_pos"].view(np.uint32))
                                               and np.array_equal(gpot, beam["potted"])
                                               and np.array_equal(gpos.view(np.uint32),
                                                                  want["final_pos"][order].view(np.uint32)))
        # ---- CPU baseline: the reference's own code over one GPU's shard (all 65 536 candidates), all host threads ----
        reference_pass(oracle, world, threads, shard=0)
        reps, tsum, evs, nst = 3, 0.0, 0, 0
        for _ in range(reps):
            out, dt = reference_pass(oracle, world, threads, shard=0)
            tsum += dt; evs += int(out["n_events"].sum()); nst += len(out["cost"])
        cpu = {"value": nst / tsum, "unit": UNIT, "cores": threads, "kind": kind, "events_per_sec": evs / tsum,
               "sample": f"{reps} passes over all {nst // reps} candidates of shard 0 (same grid and ids as the GPU arm), {tsum:.1f} s"}
    eng.set_prune(1)

    # ---- secondary workloads (device-timed, resident inputs, both prune modes) -----------------------------------------
    extra = None
    if not args.no_extra:
        extra = secondary_workloads(eng, W, args)
        extra.update(searches or {})
    barrier()

    # ---- roofline of the strike pipeline (init + 31 x {items, quartic, apply} + finish), timed together with CUDA events
    #      on the engine's stream; its FP64 work is the quartic kernel's ---------------------------------------------------
    k_ms = float(np.mean(kernel_ms))
    per_rank = cvec.cpu().numpy() / world
    ops = measured_op_counts()
    w_quartic = float(ops["fp64_flop_per_quartic"]) if ops else W_QUARTIC_FP64
    w_src = ("profiles/r02_level_fp_ops.json: FP64 thread-instructions of quartic_kernel (DADD + DMUL + 2 x DFMA) / quartics solved, ncu"
             if ops else "SURVEY.md 8d working figure (no ncu count committed)")
    fp64_alg = per_rank[2] * w_quartic          # the reference's quartic count
    fp64_exec = per_rank[5] * w_quartic         # quartics this build solves
    fp32_flop = (per_rank[1] * W_PAIR + per_rank[3] * W_RAIL + per_rank[4] * W_POCKET + per_rank[0] * nb * W_ADVANCE)
    peak_dp = eng.issue_rate(0)
    peak_dfma = eng.issue_rate(1)
    peak_sp = eng.issue_rate(2)
    peak_ffma = eng.issue_rate(3)
    alg_bytes = mine * (nb * 8 + 24)  # per strike: end state + cost/flags/potted/time written; inputs are O(1)
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            hbm_peak, hbm_src = float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        hbm_peak, hbm_src = 6650.0, "fallback (B200_PROFILING.md)"
    # DRAM traffic and warp instructions of one level's kernels from the committed ncu launch list (this workload, one GPU)
    traffic, traffic_src, warp_inst = None, None, None
    for name in ("r02_level_dram_inst_time.json", "r01_v8_level_dram_inst_time.json"):
        try:
            with open(os.path.join(ROOT, "profiles", name)) as f:
                dj = json.load(f)
            traffic = float(dj["dram_bytes_read"] + dj["dram_bytes_write"])
            warp_inst = float(dj.get("warp_instructions", 0)) or None
            traffic_src = f"profiles/{name} (ncu dram__bytes_read.sum + dram__bytes_write.sum over the level's launches)"
            break
        except Exception:
            continue
    sm_mhz = (clocks or {}).get("sm_mhz") or 1965.0
    issue_frac = (warp_inst / (k_ms * 1e-3) / (148 * 4 * sm_mhz * 1e6)) if warp_inst else None
    ach_alg = fp64_alg / (k_ms * 1e-3) / 1e9
    ach_exec = fp64_exec / (k_ms * 1e-3) / 1e9
    roofline = {"bound": "fp64_vector", "kernel": "strike pipeline (init + 31 x {items, quartic, apply} + finish; FP64 work = quartic_kernel)",
                "achieved": ach_exec, "peak": peak_dp, "unit": "GFLOP/s", "frac": ach_exec / peak_dp,
                "frac_executed": ach_exec / peak_dp,
                "frac_algorithmic": ach_alg / peak_dp,
                "frac_issue_slots": issue_frac,
                "fractions_note": "executed = quartics this build solves x measured FP64 FLOP per quartic; algorithmic = the "
                                  "reference's quartic count x the same (the reach bound skips the difference); issue slots = "
                                  "warp instructions of a level / (kernel ms x 592 issue slots per clock x SM clock)",
                "limiter_note": "a 65 536-strike level is bound by its chain of 31 dependent rounds (a 256-strike level takes 60 % of "
                                "its time), not by a pipe; at scale (extra.config5_slice) the path is issue-bound and the quartic "
                                "solver was bound by instruction fetch until quartic_kernel_wide (sm__icc hit rate 73 -> 91 %, "
                                "profiles/r02_big_round_instruction_fetch.txt; DESIGN.md section 4)",
                "fp64_flop_per_quartic": w_quartic, "fp64_flop_per_quartic_source": w_src,
                "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": "measured in this run: un-fused DADD+DMUL issue rate (FMA contraction is forbidden on "
                               "this path); DFMA rate x2 listed for context",
                "kernel_ms": k_ms, "kernel_share_of_step": k_ms / (t_ms / args.steps),
                "alg_fp64_flop_per_launch": fp64_alg, "exec_fp64_flop_per_launch": fp64_exec,
                "alg_fp32_flop_per_launch": fp32_flop, "warp_instructions_per_launch": warp_inst,
                "fp32": {"achieved": fp32_flop / (k_ms * 1e-3) / 1e9, "peak": peak_sp, "unit": "GFLOP/s"},
                "dfma_peak_gflops": 2 * peak_dfma, "ffma_peak_gflops": 2 * peak_ffma,
                "hbm": {"achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "peak_source": hbm_src},
                "work_per_strike": {"events": per_rank[0] / mine, "quartics": per_rank[2] / mine,
                                    "quartics_solved": per_rank[5] / mine},
                "exact": {"kernel_ms": float(np.mean(exact_kernel_ms)),
                          "achieved": fp64_alg / (float(np.mean(exact_kernel_ms)) * 1e-3) / 1e9,
                          "frac": fp64_alg / (float(np.mean(exact_kernel_ms)) * 1e-3) / 1e9 / peak_dp,
                          "note": "reach bound off: executed == algorithmic"}}

    value = total * args.steps / (t_ms * 1e-3)
    lm = np.array(long_ms)
    cfg = workload_desc(world)
    cfg["prune"] = ("value / e2e: reach bound ON (bcu_set_prune 1, opt-in; audited, not proven: DESIGN.md 4); value_exact / "
                    "e2e_exact: OFF (the C ABI's default)")
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "untimed_steps_before_timed_region": max(args.warmup, 3) + SETTLE,
            "ms_per_step": t_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32+f64", "data": "synthetic",
            "config": cfg, "events_per_sec": ev_total * args.steps / (t_ms * 1e-3),
            "value_exact": total * args.steps / (t_exact_ms * 1e-3), "ms_per_step_exact": t_exact_ms / args.steps,
            "wall_s_timed_region": wall, "clocks": clocks,
            "stats": {"steps": long_steps, "value": total * long_steps / (long_sum * 1e-3), "ms_mean": float(lm.mean()),
                      "ms_median": float(np.median(lm)), "ms_p95": float(np.percentile(lm, 95)), "ms_min": float(lm.min()),
                      "ms_max": float(lm.max()), "note": "rank 0's per-step device times; value = max-over-ranks sum"},
            "e2e": {"value": total * e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "steps": e2e_steps},
            "e2e_exact": {"value": total * e2e_steps / e2e_exact_s, "unit": UNIT, "steps": e2e_steps},
            "gpu_launches": launches, "parity": parity, "roofline": roofline, "cpu_baseline": cpu,
            "best_strike": {"cost": int(beam["cost"][0]), "id": int(beam["id"][0])} if len(beam["cost"]) else None,
            "extra": extra}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def secondary_workloads(eng, W, args):
    """Device-timed secondary measurements on rank 0's GPU (resident inputs, L2 flushed before each timed launch, CUDA
    events on the engine's stream; both prune modes): the 16-ball level the north_star target is quoted on and a slice of
    config 5 (random 16-ball tables x 4 096 strikes)."""
    out = {}
    red = 1 << W.TYPE_CODE["RED"]

    def timed(n_strikes, reps):
        res = {}
        for prune, key in ((1, "strikes_per_sec"), (0, "strikes_per_sec_exact")):
            eng.set_prune(prune)
            eng.rerun(); eng.rerun()
            ms = []
            for _ in range(reps):
                eng.flush_l2()
                eng.timer_begin()
                eng.rerun()
                ms.append(eng.timer_end())
            res[key] = n_strikes / (float(np.mean(ms)) * 1e-3)
            res["ms" if prune else "ms_exact"] = float(np.mean(ms))
        eng.enable_counters(True)
        eng.rerun()
        res["events_per_sec_exact"] = eng.counters()["events"] / (res["ms_exact"] * 1e-3)
        eng.enable_counters(False)
        return res
    pos, types = W.random_layout(16, 16)
    dirs, speeds = W.grid(N_DIRS, N_SPEEDS)
    eng.set_states(pos, types, red)
    eng.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST)
    out["level_16_balls"] = dict(workload="16-ball layout (seed 16), 1024 x 64 = 65536 strikes, one level", **timed(65536, 10))
    nt = 1000
    tp, tt = W.random_tables(nt, 16, seed0=1000)
    d5, s5 = W.grid(64, 64)
    eng.set_states(tp, tt, red)
    eng.evaluate_grid(d5, s5, W.GRID_SEARCH_COST)
    out["config5_slice"] = dict(workload=f"config 5: tables 1000..{1000 + nt - 1} (16 balls) x 4096 strikes = {nt * 4096} strikes, one launch",
                                **timed(nt * 4096, 3))
    eng.set_prune(1)
    return out


def search_workloads(eng, W, rank, world, dist):
    """BASELINE configs 4 and 3 through the C ABI's bcu_search_levels (the whole level loop in C++: evaluate -> device top-k ->
    allgather + merge -> rules): per-level device time, wall time of the whole search (of the bcu_search_levels call itself; the
    time including this binding's conversion of 1024 solutions into Python objects is listed beside it), a digest of the
    solution list (the same for every N), both prune modes.  Config 4 is STRONG scaling: ~1 M candidates per level sharded over the N ranks."""
    import hashlib
    from billiard_ai_b200.rules import RepeatRules, SnookerRules
    from billiard_ai_b200.search import search_levels
    out = {}

    def run(name, desc, pos, types, levels, beam, dirs, speeds, rules):
        res = {"workload": desc, "n_gpus": world}
        for prune, tag in ((1, ""), (0, "_exact")):
            eng.set_prune(prune)
            best = None
            for it in range(3):   # the first pass warms up (allocations, graph capture)
                stats, info = [], {}
                t0 = time.perf_counter()
                sols = search_levels(eng, pos, types, [W.TYPE_CODE["RED"]], levels, beam, dirs, speeds, rules, stats=stats, info=info)
                wall = time.perf_counter() - t0
                if it > 0 and (best is None or info["wall_ms"] < best[0]):
                    best = (info["wall_ms"], stats, sols, wall)
            wall, stats, sols, wall_py = best[0] * 1e-3, best[1], best[2], best[3]   # wall = the C-ABI call itself
            digest = hashlib.sha256(json.dumps([(s["cost"], s["dirs"], s["speeds"]) for s in sols]).encode()).hexdigest()[:16]
            if dist is not None:
                allv = [None] * world
                dist.all_gather_object(allv, (digest, wall, [st["kernel_ms"] for st in stats], wall_py))
                assert len({v[0] for v in allv}) == 1, "ranks disagree on the solution list"
                wall = max(v[1] for v in allv)
                wall_py = max(v[3] for v in allv)
                kms = [max(v[2][i] for v in allv) for i in range(len(stats))]
            else:
                kms = [st["kernel_ms"] for st in stats]
            strikes = sum(st["candidates"] for st in stats)
            res.update({"strikes": strikes, "solutions": len(sols), "digest" + tag: digest,
                        "best" + tag: [(s["cost"], s["breaks"]) for s in sols[:3]],
                        "wall_ms" + tag: 1e3 * wall, "wall_ms_incl_python_result_objects" + tag: 1e3 * wall_py,
                        "device_ms" + tag: float(sum(kms)),
                        "level_kernel_ms" + tag: kms, "level_candidates": [st["candidates"] for st in stats],
                        "strikes_per_sec_wall" + tag: strikes / wall, "wall_over_device" + tag: 1e3 * wall / float(sum(kms))})
        assert res["digest"] == res["digest_exact"], "reach bound on / off find different solutions"
        out[name] = res

    if world == 1:
        out["planner"] = planner_workload(eng, W)
    pos, types = W.snooker22_open()
    d4, s4 = W.grid(32, 32)
    run("config4", "config 4: snooker 22-ball frame, 4-break search, beam 1024 x (32 directions x 32 speeds) = 1 048 576 strikes per "
        f"level from level 3 on, sharded over {world} GPU(s), snooker rules between breaks; times = max over ranks",
        pos, types, 4, 1024, d4, s4, SnookerRules())
    if world == 1:
        p3, t3 = W.random_layout(16, 16)
        d3, s3 = W.grid(256, 16)
        run("config3", "config 3: pool 16-ball layout (seed 16), 3-break beam search, beam 1024 x (256 directions x 16 speeds) = "
            "4 194 304 strikes in level 3, 'repeat' rules", p3, t3, 3, 1024, d3, s3, RepeatRules())
    return out


def planner_workload(eng, W):
    """The reference's geometric planner on the device (bcu_plan) + the simulation of what it plans (bcu_evaluate_plan): config 1
    (the reference's own test layout, one state) and a beam of 1024 random 16-ball states planned in one call.  Device time by
    CUDA events on the engine's stream; the reference plans one state at a time on the host (1 700 expansions, ~47 ms)."""
    import json as _json
    tbl = W.default_table()
    eng.set_planner(tbl, velocity_multiply=True)
    red = 1 << W.TYPE_CODE["RED"]
    out = {}
    with open(os.path.join(ROOT, "tests", "golden", "layouts.json")) as f:
        lay = _json.load(f)["test_01"]
    p1 = np.array([b["position"] for b in lay["balls"]], np.float32)
    t1 = np.array([W.TYPE_CODE[b["type"]] for b in lay["balls"]], np.int32)
    tp, tt = W.random_tables(1024, 16, seed0=5000)
    for name, pos, types, desc in (("config1", p1[None], t1[None], "config 1: 8-ball layout 00_test_01 of test/billiard_search, single-break search, "
                                    "candidates = the reference planner's own (direct + up to 3 cushions, 3 speeds)"),
                                   ("beam1024", tp, tt, "1024 random 16-ball states (seeds 5000..6023) planned and simulated in one call each")):
        eng.set_states(pos, types, red)
        eng.set_prune(0)
        n = eng.plan()
        res = eng.evaluate_plan(final_pos=False)
        plan_ms, eval_ms = [], []
        for _ in range(5):
            eng.timer_begin(); eng.plan(); plan_ms.append(eng.timer_end())
            eng.timer_begin(); eng.evaluate_plan(final_pos=False); eval_ms.append(eng.timer_end())
        valid = int(((res["flags"] & 2) != 0).sum())
        out[name] = dict(workload=desc, states=int(len(pos)), strikes_planned=int(n), valid_strikes=valid,
                         best_cost=int(res["cost"][(res["flags"] & 2) != 0].min()) if valid else None,
                         plan_ms=float(np.median(plan_ms)), evaluate_ms=float(np.median(eval_ms)),
                         states_per_sec=len(pos) / ((np.median(plan_ms) + np.median(eval_ms)) * 1e-3))
    eng.set_prune(1)
    return out


def emit(line):
    """The ONE JSON line goes to the real stdout; everything else (NCCL banners, library chatter) was sent to stderr."""
    REAL_STDOUT.write(json.dumps(line) + "\n")
    REAL_STDOUT.flush()


REAL_STDOUT = sys.stdout


def main():
    global REAL_STDOUT
    REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)  # native libraries that print to fd 1 (e.g. "NCCL version ...") must not pollute the result line
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the reference passes (parity + cpu_baseline)")
    ap.add_argument("--no-extra", action="store_true", help="skip the secondary workloads")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("bench.py --gpus N>1 must be launched with torch.distributed.run --nproc-per-node N")
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_b200(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
