#!/usr/bin/env python
"""bench.py -- strikes/s of the strike-evaluation hot path (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" is one break level of the search over one batch of synthetic candidates (BASELINE.json configs[1],
restated in SURVEY.md 8d "Config 2"): the 22-ball snooker layout x (1024 directions x 64 speeds) = 65 536
candidate strikes PER GPU (weak scaling: N GPUs evaluate N x 1024 directions, sharded by contiguous id range);
each strike is simulated to rest, costed, the k = 1024 best valid strikes are selected on the device and -- for
N > 1 -- merged across GPUs with one ncclAllGather of the per-GPU record blocks.

  value : whole-job strikes/s, inputs resident in HBM (CUDA events on the engine's stream, max over ranks)
  e2e   : the same level through the host-buffer C-ABI calls (bcu_set_states + bcu_evaluate_level (= bcu_evaluate + bcu_level_select) +
          bcu_fetch_beam): pinned host candidate arrays in; per-candidate cost / flags / potted mask / final time and the
          beam (k best strikes with their final table positions) out, wall clock
  roofline     : the simulate kernel's algorithmic FP64 work / its CUDA-event time vs. the FP64 vector issue peak
                 measured in the same run (un-fused DADD+DMUL: FMA contraction is forbidden on this path)
  cpu_baseline : the reference's own CPU code (oracle/_ref/libref_oracle.so) on a bounded sample, all host threads

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


def reference_sample(world, n_sample):
    from billiard_ai_b200 import workloads as W
    dirs, speeds = sharded_grid(world)
    vx, vy = W.grid_velocities(dirs, speeds)
    total = len(vx)
    idx = (np.arange(n_sample, dtype=np.int64) * 2654435761) % total  # fixed pseudo-random sample of the grid
    return vx[idx], vy[idx]


def time_reference(oracle, world, n_sample, threads):
    from billiard_ai_b200 import workloads as W
    pos, types = W.snooker22_open()
    vx, vy = reference_sample(world, n_sample)
    t0 = time.perf_counter()
    out = oracle.evaluate(pos[None], types[None], 1 << W.TYPE_CODE["RED"], 0, vx, vy, W.GRID_SEARCH_COST, 0,
                          threads=threads)
    dt = time.perf_counter() - t0
    return n_sample / dt, int(out["n_events"].sum()) / dt, dt


def calibrate_sample(oracle, world, threads, target_s):
    n = 64 * threads
    for _ in range(2):  # the first probe is dominated by thread start-up
        rate, _, _ = time_reference(oracle, world, n, threads)
        n = (int(max(256, min(1 << 20, rate * target_s))) + 63) // 64 * 64
    return n


def run_reference(args, rank, world):
    if rank != 0:
        return
    oracle, kind = load_reference()
    threads = os.cpu_count() or 1
    n_sample = calibrate_sample(oracle, world, threads, target_s=4.0)
    for _ in range(min(args.warmup, 1)):
        time_reference(oracle, world, max(256, n_sample // 8), threads)
    tot_t, tot_n, tot_ev = 0.0, 0, 0.0
    for _ in range(args.steps):
        r, ev, dt = time_reference(oracle, world, n_sample, threads)
        tot_t += dt; tot_n += n_sample; tot_ev += ev * dt
    value = tot_n / tot_t
    sample = f"{n_sample} of {N_DIRS * N_SPEEDS * world} grid candidates per step (fixed multiplicative-hash sample)"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32+f64", "data": "synthetic",
            "config": workload_desc(world), "events_per_sec": tot_ev / tot_t,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ---- B200 arm -------------------------------------------------------------------------------------------------
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
    dirs, speeds = sharded_grid(world)
    total = len(dirs) * len(speeds)
    lo, hi = W.shard_range(total, rank, world)
    mine = hi - lo
    eng.set_states(pos, types, 1 << W.TYPE_CODE["RED"])
    eng.evaluate_grid(dirs, speeds, W.GRID_SEARCH_COST, first=lo, count=mine)   # uploads + first launch

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        eng.sync()

    # ---- warm-up ------------------------------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        eng.level_rerun(BEAM)
    eng.sync()

    # ---- timed: value (resident inputs) ----------------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    launches0 = eng.launch_count()
    barrier()
    wall0 = time.perf_counter()
    step_ms, kernel_ms = [], []
    for _ in range(args.steps):
        eng.flush_l2()
        eng.timer_begin()
        eng.level_rerun(BEAM)
        step_ms.append(eng.timer_end())
        kernel_ms.append(eng.last_kernel_ms())
    barrier()
    wall = time.perf_counter() - wall0
    launches = eng.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    t_ms = float(np.sum(step_ms))
    if world > 1:
        t = torch.tensor([t_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        t_ms = float(t.item())
    beam = eng.fetch_beam()

    # ---- work counters (untimed extra launch) -> events/s and the roofline's algorithmic work ------------------
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

    # ---- e2e: host buffers through the C ABI -----------------------------------------------------------------------
    vx_all, vy_all = W.grid_velocities(dirs, speeds)

    def pinned(a):
        return torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
    keep = [pinned(vx_all[lo:hi]), pinned(vy_all[lo:hi]), pinned(np.zeros(mine, np.int32)),
            pinned(np.full(mine, W.GRID_SEARCH_COST, np.int64)), pinned(np.zeros(mine, np.int64))]
    hvx, hvy, hst, hsc, hbc = [t.numpy() for t in keep]
    hsc = hsc.view(np.uint64); hbc = hbc.view(np.uint64)
    eng.set_id_base(lo)

    def e2e_step():
        eng.set_states(pos, types, 1 << W.TYPE_CODE["RED"])
        res = eng.evaluate(hst, hvx, hvy, hsc, hbc, final_pos=False, event_log=False, reuse_buffers=True, select_k=BEAM)
        b = eng.fetch_beam()
        return res, b
    e2e_steps = max(1, min(args.steps, 10))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        res, b2 = e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    assert np.array_equal(b2["cost"], beam["cost"]) and np.array_equal(b2["id"], beam["id"]), "e2e beam != resident beam"
    nb = len(types)
    h2d = mine * (4 + 4 + 4 + 8 + 8) + nb * (8 + 4) + 4
    d2h = mine * (8 + 4 + 4 + 4) + BEAM * (24 + nb * 8)   # per-strike cost / flags / potted / time + the beam records

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the simulate stage: the strike pipeline's kernels (items / quartic / apply per event round), timed
    #      together with CUDA events on the engine's stream; its FP64 work is the quartic kernel's ------------------------
    k_ms = float(np.mean(kernel_ms))
    per_rank = cvec.cpu().numpy() / world
    fp64_flop = per_rank[2] * W_QUARTIC_FP64
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
    # DRAM traffic of one level's kernels from the committed ncu launch list (warm caches, single GPU, this workload)
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "r01_v8_level_dram_inst_time.json")) as f:
            dj = json.load(f)
        traffic = float(dj["dram_bytes_read"] + dj["dram_bytes_write"])
        traffic_src = "profiles/r01_v8_level_dram_inst_time.json (ncu dram__bytes_read.sum + dram__bytes_write.sum over the level's launches)"
    except Exception:
        pass
    ach = fp64_flop / (k_ms * 1e-3) / 1e9
    roofline = {"bound": "fp64_vector", "kernel": "strike pipeline (init + 31 x {items, quartic, apply} + finish; FP64 work = quartic_kernel)", "achieved": ach, "peak": peak_dp,
                "unit": "GFLOP/s", "frac": ach / peak_dp, "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": "measured in this run: un-fused DADD+DMUL issue rate (FMA contraction is forbidden on "
                               "this path); DFMA rate x2 listed for context",
                "kernel_ms": k_ms, "kernel_share_of_step": k_ms / (t_ms / args.steps),
                "alg_fp64_flop_per_launch": fp64_flop, "alg_fp32_flop_per_launch": fp32_flop,
                "fp32": {"achieved": fp32_flop / (k_ms * 1e-3) / 1e9, "peak": peak_sp, "unit": "GFLOP/s"},
                "dfma_peak_gflops": 2 * peak_dfma, "ffma_peak_gflops": 2 * peak_ffma,
                "hbm": {"achieved": alg_bytes / (k_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "peak_source": hbm_src},
                "work_per_strike": {"events": per_rank[0] / mine, "quartics": per_rank[2] / mine,
                                    "quartics_solved": per_rank[5] / mine},
                "work_note": "algorithmic work = what the reference computes for these strikes (SURVEY.md 8d); the reach "
                             "bound (DESIGN.md 4) leaves quartics_solved of the quartics per strike to the FP64 pipe"}

    # ---- CPU baseline (reference's own code, bounded sample) ----------------------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        oracle, kind = load_reference()
        threads = os.cpu_count() or 1
        n_sample = calibrate_sample(oracle, 1, threads, target_s=8.0)
        rate, ev_rate, dt = time_reference(oracle, 1, n_sample, threads)
        cpu = {"value": rate, "unit": UNIT, "cores": threads, "kind": kind, "events_per_sec": ev_rate,
               "sample": f"{n_sample} of {total} grid candidates (fixed multiplicative-hash sample), {dt:.1f} s wall"}

    value = total * args.steps / (t_ms * 1e-3)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": t_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32+f64", "data": "synthetic",
            "config": workload_desc(world), "events_per_sec": ev_total * args.steps / (t_ms * 1e-3),
            "wall_s_timed_region": wall, "clocks": clocks,
            "e2e": {"value": total * e2e_steps / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "steps": e2e_steps},
            "gpu_launches": launches, "roofline": roofline, "cpu_baseline": cpu,
            "best_strike": {"cost": int(beam["cost"][0]), "id": int(beam["id"][0])} if len(beam["cost"]) else None}
    emit(line)
    if world > 1:
        dist.destroy_process_group()


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
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
