// K1 v3: the event-synchronous strike pipeline (the batch path of bcu_evaluate / bcu_evaluate_grid).
//
// Same arithmetic as simulate_candidates (bcu_kernels.cu, kept for the single-strike trace), but decomposed so that
// every kernel is small (fits the instruction caches), dense (one thread per *item*, not per ball) and can be tuned
// on its own.  ncu on the warp-per-strike kernels showed what bounds this path: 13-15 of 32 lanes active and an
// instruction-fetch stall (`no_instruction`) of 24 cycles per issue from a 160 KB instruction footprint.
//
// All strikes of a launch advance one event per ROUND (the reference's event loop, billiard_search.cpp:1370-1381,
// turned inside out).  State lives in HBM/L2 as structure-of-arrays; per round three kernels run on the stream:
//
//   items_kernel    one thread per dynamic ball of every live strike walks its rails, pockets and partner balls
//                   (nextRailCollision :1639-1682, nextPocketCollision :1684-1714, nextBallCollision :1603-1637)
//                   with the cheap first-stage tests; survivors are compacted per CTA in shared memory so the second
//                   stages (rail / pocket quadratics, quantised quartic coefficients) run on dense threads.
//                   Rail / pocket times go to the strike's 64-bit (time, evaluation order) key with atomicMin -- the
//                   reference's strict-'<' first-evaluated-wins choice (:2002-2013); ball pairs become quartic jobs
//                   in a global queue, filled from both ends by the arm of the solver a job will take.
//   quartic_kernel  one thread per job: the FP64 complex closed-form quartic (billiard_physics.cpp:470-560).  Large parts
//                   take quartic_kernel_wide: CTAs whose warps walk the solver side by side (instruction cache).
//   apply_kernel    one thread per live strike applies the winning event (createLayer :1957-1992, System::append
//                   type.cpp:467-531) and one thread per other dynamic ball advances it in closed form; both
//                   offer the rest / rolling times of the next round (nextBallInRest :1590, nextRolling :1575) and
//                   build the next round's stream-compacted lists of live strikes and dynamic balls.
//
// After the last round finish_kernel (one warp per strike) evaluates the rules and simulationCost (:1388-1405, :1473-1532)
// and writes the results.  Keys are triple-buffered by round (read r%3, offered (r+1)%3, reset (r+2)%3) and layer times double-buffered, so no
// kernel reads a word another thread of the same launch writes.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "bcu_kernels.cuh"
#include "bcu_physics.cuh"

namespace bcu {

namespace {

__constant__ PhysConst p_phys;

constexpr int ROUNDS = MAX_EVENTS + 1;  // round r applies event r+1; round MAX_EVENTS only flags overflow
constexpr int QUARTIC_BLOCK = 128;
#ifndef QUARTIC_MIN_BLOCKS
#define QUARTIC_MIN_BLOCKS 6
#endif
#ifndef QUARTIC_WIDE_BLOCK_V
#define QUARTIC_WIDE_BLOCK_V 512      // quartic_kernel_wide: the kernel for large parts (see launch_pipeline)
#endif
#ifndef QUARTIC_WIDE_MIN_BLOCKS
#define QUARTIC_WIDE_MIN_BLOCKS 2
#endif
constexpr int QUARTIC_WIDE_BLOCK = QUARTIC_WIDE_BLOCK_V;
constexpr int APPLY_BLOCK = 256;

enum Category : uint32_t { CAT_REST = 0, CAT_ROLLING = 1, CAT_STATIC_PAIR = 2, CAT_DYNAMIC_PAIR = 3, CAT_RAIL = 4, CAT_POCKET = 5 };

__device__ __forceinline__ uint32_t ordered_bits(float t) {
    const uint32_t u = __float_as_uint(t + 0.0f);  // -0 -> +0
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float from_ordered_bits(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}
__device__ __forceinline__ uint32_t make_code(uint32_t ball, uint32_t cat, uint32_t partner) {
    return (ball << 16) | (cat << 8) | partner;
}
__device__ __forceinline__ void offer(unsigned long long* key, float t, uint32_t code) {
    if (t != t) return;  // a NaN time never wins the reference's strict '<'
    atomicMin(key, ((unsigned long long)ordered_bits(t) << 32) | code);
}

// per-ball meta word: kind | rolling << 8 | (lastRail + 1) << 16 | (type + 1) << 24
__device__ __forceinline__ uint32_t make_meta(int kind, bool rolling, int lastRail, int type) {
    return (uint32_t)kind | ((uint32_t)rolling << 8) | ((uint32_t)(lastRail + 1) << 16) | ((uint32_t)(type + 1) << 24);
}
__device__ __forceinline__ int meta_kind(uint32_t m) { return (int)(m & 0xffu); }
__device__ __forceinline__ bool meta_rolling(uint32_t m) { return ((m >> 8) & 1u) != 0; }
__device__ __forceinline__ int meta_last_rail(uint32_t m) { return (int)((m >> 16) & 0xffu) - 1; }
__device__ __forceinline__ int meta_type(uint32_t m) { return (int)(m >> 24) - 1; }
__device__ __forceinline__ bool kind_dynamic(int kind) {
    return kind == K_SHOT || kind == K_MOVING || kind == K_COLLISION || kind == K_RAIL;
}

// acceleration of a ball at rest: the BallState ctor's accelerationLength * normalize(0) = (-0, -0) (type.cpp:176-184)
__device__ __forceinline__ v2 rest_acc(const PhysConst& pc) { return pc.slideAccLen * mk(0.0f, 0.0f); }

struct Candidate {
    int sIdx;
    float vx, vy;
    uint64_t searchCost, baseCost;
};
__device__ __forceinline__ Candidate decode_candidate(const LaunchParams& p, int64_t cand) {
    Candidate c;
    if (p.src.gridDirs > 0) {
        const int64_t gid = cand + p.src.firstCandidate;
        const int64_t perState = (int64_t)p.src.gridDirs * p.src.gridSpeeds;
        c.sIdx = (int)(gid / perState);
        const int rem = (int)(gid - (int64_t)c.sIdx * perState);
        const int k = rem / p.src.gridSpeeds, j = rem - k * p.src.gridSpeeds;
        const float sp = p.src.speedTable[j];
        c.vx = sp * p.src.dirTable[2 * k];
        c.vy = sp * p.src.dirTable[2 * k + 1];
        c.searchCost = p.src.gridSearchCost;
        c.baseCost = p.src.stateBaseCost ? p.src.stateBaseCost[c.sIdx] : 0ull;
    } else {
        c.sIdx = p.src.stateIdx[cand];
        c.vx = p.src.vx[cand];
        c.vy = p.src.vy[cand];
        c.searchCost = p.src.searchCost[cand];
        c.baseCost = p.src.baseCost[cand];
    }
    return c;
}

// Reserves list slots for up to three predicates per thread with ONE global atomicAdd per predicate and CTA (same-address
// atomics are the serial part of stream compaction).  All threads of the CTA must call; *slotK is valid where predK.
struct AppendShared {
    uint32_t warpCount[4][8];
    uint32_t warpBase[4][8];
};
template <int NPRED>
__device__ __forceinline__ void cta_reserve(AppendShared& sh, const bool (&pred)[NPRED], uint32_t* const (&counter)[NPRED],
                                            uint32_t (&slot)[NPRED]) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nWarps = blockDim.x >> 5;
    unsigned m[NPRED];
    #pragma unroll
    for (int k = 0; k < NPRED; ++k) {
        m[k] = __ballot_sync(FULL, pred[k]);
        if (lane == 0) sh.warpCount[k][warp] = __popc(m[k]);
    }
    __syncthreads();
    if (threadIdx.x < NPRED) {
        const int k = threadIdx.x;
        uint32_t total = 0;
        for (int i = 0; i < nWarps; ++i) { sh.warpBase[k][i] = total; total += sh.warpCount[k][i]; }
        const uint32_t base = total ? atomicAdd(counter[k], total) : 0u;
        for (int i = 0; i < nWarps; ++i) sh.warpBase[k][i] += base;
    }
    __syncthreads();
    #pragma unroll
    for (int k = 0; k < NPRED; ++k) slot[k] = sh.warpBase[k][warp] + __popc(m[k] & ((1u << lane) - 1u));
}

// A strike that reached its last layer is only MARKED here (flags | nEvents << 8, final layer time); the rules, the cost
// and the result write-out are done for all strikes at once by finish_kernel.
__device__ __forceinline__ void mark_finished(const LaunchParams& p, int64_t s, uint32_t flags, int nEvents, float T) {
    p.out.flags[s] = flags | ((uint32_t)nEvents << 8);
    p.out.finalTime[s] = T;
}

// ---------------------------------------------------------------------------------------------------------------------
// init: toInputLayer (billiard_search.cpp:760-795) + System::append of layer 0.  One thread per strike (the per-ball
// arrays are ball-major, [ball][strike], so consecutive threads write consecutive words).
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void init_body(AppendShared& ash, const LaunchParams& p, const PipeWorkspace& w, unsigned bid, unsigned nb) {
    const int NB = p.states.nBalls;
    const PhysConst pc = p_phys;
    const int64_t n = w.n;
    const int BS = blockDim.x;
    for (int64_t base = (int64_t)bid * BS; base < p.nCand; base += (int64_t)nb * BS) {
        const int64_t s = base + threadIdx.x;
        uint32_t dynMask = 0, restMask = 0;
        if (s < p.nCand) {
            const Candidate c = decode_candidate(p, s);
            unsigned long long* key = w.key + s;  // buffer 0
            for (int b = 0; b < NB; ++b) {
                const int type = p.states.type[(int64_t)c.sIdx * NB + b];
                int kind = K_ABSENT;
                v2 pos = mk(0.f, 0.f), vel = mk(0.f, 0.f), vn, acc;
                float nextRolling = -1.0f;
                if (type >= 0) {
                    const float2 q = reinterpret_cast<const float2*>(p.states.pos)[(int64_t)c.sIdx * NB + b];
                    pos = mk(q.x, q.y);
                    if (type == p.rules->cueType) {
                        vel = mk(c.vx, c.vy);
                        kind = K_SHOT;
                        nextRolling = time_to_rolling(vel, pc);  // type.cpp:502-506
                        dynMask |= 1u << b;
                    } else {
                        kind = K_REST;
                        restMask |= 1u << b;
                    }
                }
                const int64_t g = (int64_t)b * n + s;
                w.pos[g] = make_float2(pos.x, pos.y);
                w.meta[g] = make_meta(kind, false, -1, type);
                if (w.merged) {   // balls that do not move are valid in both layer buffers
                    w.buf[1].pos[g] = make_float2(pos.x, pos.y);
                    w.buf[1].meta[g] = make_meta(kind, false, -1, type);
                }
                // A ball at rest is position + meta: its velocity (0, 0), normalised velocity (0, 0), acceleration (-0, -0) and
                // "no rolling time" are constants of the BallState ctor (rest_ball()), so they are neither written here nor read
                // back (880 -> 292 bytes per 22-ball strike: the level's DRAM traffic was 6 x its algorithmic bytes)
                if (kind == K_SHOT) {
                    derive_ball(vel, false, pc, &vn, &acc);
                    w.vel[g] = make_float2(vel.x, vel.y);
                    w.vn[g] = make_float2(vn.x, vn.y);
                    w.acc[g] = make_float2(acc.x, acc.y);
                    w.nextRolling[g] = nextRolling;
                }
                if (kind == K_SHOT) {
                    offer(key, time_to_stop(acc, vel), make_code(b, CAT_REST, 0));
                    if (nextRolling >= 0.0f) offer(key, nextRolling, make_code(b, CAT_ROLLING, 0));
                }
            }
            w.dynMask[s] = dynMask; w.restMask[s] = restMask; w.T[s] = 0.0f;
            if (w.merged) w.stopped[0][s] = 0xffu;
            if (dynMask == 0) mark_finished(p, s, F_COMPLETED, 0, 0.0f);
        }
        // live strikes, and their dynamic balls (normally the one cue ball; further cue-type balls are appended one by one)
        const bool pred[2] = {dynMask != 0, dynMask != 0};
        uint32_t* const counter[2] = {&w.counts[0], &w.counts[1]};
        uint32_t slot[2];
        cta_reserve<2>(ash, pred, counter, slot);
        if (dynMask != 0) {
            w.live[slot[0]] = (uint32_t)s;
            w.dyn[slot[1]] = ((uint32_t)s << 5) | (uint32_t)(__ffs(dynMask) - 1);
            for (uint32_t m = dynMask & (dynMask - 1); m; m &= m - 1)
                w.dyn[atomicAdd(&w.counts[1], 1u)] = ((uint32_t)s << 5) | (uint32_t)(__ffs(m) - 1);
        }
    }
}

__global__ void __launch_bounds__(256) init_kernel(const LaunchParams* __restrict__ pp, PipeWorkspace w) {
    if (w.pdlEarly) cudaTriggerProgrammaticLaunchCompletion();   // see launch_pipeline
    __shared__ AppendShared ash;
    init_body(ash, *pp, w, blockIdx.x, gridDim.x);
}

#include "bcu_items.cuh"

// ---------------------------------------------------------------------------------------------------------------------
// quartic: one thread per job.
// ---------------------------------------------------------------------------------------------------------------------
// The job queue is filled from both ends by items_kernel (bcu_items.cuh stage 2): "real" jobs -- the solver's std::pow(complex,
// double) takes its real arm -- from slot 0 upwards, "complex" jobs from the last slot downwards.  A job is its five coefficients
// and its code, and its result reaches the strike's key through an atomicMin, so it does not matter which thread solves it.
struct JobQueue {
    uint64_t nFront, nBack;   // jobs in [0, nFront) and in (jobCap - 1 - nBack, jobCap - 1]
};
__device__ __forceinline__ JobQueue job_queue(const PipeWorkspace& w, int round) {
    const uint32_t postedFront = w.counts[4 * round + 2], postedBack = w.counts[4 * round + 3];
    JobQueue q;
    q.nFront = postedFront < w.jobCap ? postedFront : w.jobCap;
    q.nBack = postedBack < w.jobCap - q.nFront ? postedBack : w.jobCap - q.nFront;
    if ((uint64_t)postedFront + postedBack > w.jobCap && blockIdx.x == 0 && threadIdx.x == 0)
        w.counts[4 * ROUNDS + 3] = 1u;   // the two ends met: the host grows the queue and re-runs the launch
    return q;
}
__device__ __forceinline__ void quartic_solve_job(const PipeWorkspace& w, unsigned long long* key, uint64_t j) {
    float co[5];
    #pragma unroll
    for (int q = 0; q < 5; ++q) co[q] = w.jobCo[(size_t)q * w.jobCap + j];
    const uint32_t c2 = w.jobCode[j];
    float t;
    if (collision_time_from_coefficients(co, &t)) {
        const uint32_t ent = (c2 & ~JOB_MARK) >> 6;
        if (c2 & JOB_MARK) {
            // audit (pruneMode 2): the reach bound would have skipped this job, so its time must lose against the
            // strike's best so far (which only decreases); anything else is a failure of the bound
            const float tNow = from_ordered_bits((uint32_t)(__ldcg(key + (ent >> 5)) >> 32));
            if (!(t > tNow)) atomicAdd(&w.counts[4 * ROUNDS + 2], 1u);
        }
        offer(key + (ent >> 5), t, make_code(ent & 31u, (c2 & 32u) ? CAT_DYNAMIC_PAIR : CAT_STATIC_PAIR, c2 & 31u));
    }
}

// Small parts (below PipeLaunchCfg::denseMinQuartic strikes): 128 threads x 6 CTAs per SM, 80 registers, one job per thread
// and trip -- the rounds of such a part are latency-bound and want fast threads.
__global__ void __launch_bounds__(QUARTIC_BLOCK, QUARTIC_MIN_BLOCKS) quartic_kernel(PipeWorkspace w, int round) {
    if (w.pdlEarly) cudaTriggerProgrammaticLaunchCompletion();
    cudaGridDependencySynchronize();
    const JobQueue q = job_queue(w, round);
    unsigned long long* key = w.key + (w.fused ? (size_t)0 : (size_t)(round % 3) * w.n);
    const uint64_t nJobs = q.nFront + q.nBack;
    for (uint64_t v = (uint64_t)blockIdx.x * QUARTIC_BLOCK + threadIdx.x; v < nJobs; v += (uint64_t)gridDim.x * QUARTIC_BLOCK)
        quartic_solve_job(w, key, v < q.nFront ? v : w.jobCap - 1u - (v - q.nFront));
}

// Large parts: the solver is bound by INSTRUCTION FETCH when the warps of an SM are spread over it -- a job walks through 39 KB
// of code, the SM's instruction cache holds ~32 KB (sm__icc hit rate 73 %, `no instruction` the largest stall,
// profiles/r02_big_round_instruction_fetch.txt).  Here the warps of a CTA start every tile of jobs TOGETHER (a barrier per
// tile) and the tiles are of one class (front region first, then the back region; no tile straddles them), so the warps walk the
// code side by side and need a few KB of it at any time: 512 threads x 2 CTAs per SM, 64 registers.  Measured against the same
// kernel without the barrier and with 128-thread CTAs (profiles/r02_ab_experiments.txt): 1000 tables x 4096 strikes 116.6 against
// 105.1 M strikes/s (81.9 against 72.6 M with the reach bound off); parts of 2^18 strikes 120 against 110 M.
__global__ void __launch_bounds__(QUARTIC_WIDE_BLOCK, QUARTIC_WIDE_MIN_BLOCKS) quartic_kernel_wide(PipeWorkspace w, int round) {
    if (w.pdlEarly) cudaTriggerProgrammaticLaunchCompletion();
    cudaGridDependencySynchronize();
    const JobQueue q = job_queue(w, round);
    unsigned long long* key = w.key + (w.fused ? (size_t)0 : (size_t)(round % 3) * w.n);
    const uint64_t tilesFront = (q.nFront + QUARTIC_WIDE_BLOCK - 1) / QUARTIC_WIDE_BLOCK;
    const uint64_t tilesBack = (q.nBack + QUARTIC_WIDE_BLOCK - 1) / QUARTIC_WIDE_BLOCK;
    for (uint64_t t = blockIdx.x; t < tilesFront + tilesBack; t += gridDim.x) {
        __syncthreads();
        const bool front = t < tilesFront;
        const uint64_t v = (front ? t : t - tilesFront) * QUARTIC_WIDE_BLOCK + threadIdx.x;
        if (v < (front ? q.nFront : q.nBack)) quartic_solve_job(w, key, front ? v : w.jobCap - 1u - v);
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// apply: the winning event of every live strike, the closed-form advance of every other dynamic ball, next lists.
// ---------------------------------------------------------------------------------------------------------------------
struct BallRec {
    v2 pos, vel, vn, acc;
    float nextRolling;
    int kind, lastRail, type;
    bool rolling;
};
__device__ __forceinline__ BallRec load_ball(const PipeWorkspace& w, int64_t g) {
    BallRec r;
    const float2 q0 = w.pos[g], q1 = w.vel[g], q2 = w.vn[g], q3 = w.acc[g];
    const uint32_t m = w.meta[g];
    r.pos = mk(q0.x, q0.y); r.vel = mk(q1.x, q1.y); r.vn = mk(q2.x, q2.y); r.acc = mk(q3.x, q3.y);
    r.nextRolling = w.nextRolling[g];
    r.kind = meta_kind(m); r.rolling = meta_rolling(m); r.lastRail = meta_last_rail(m); r.type = meta_type(m);
    return r;
}
// A ball at rest (kind REST since init, or after a REST event): only position and meta are kept current in the workspace;
// the rest of its BallState is what the ctor derives from a zero velocity (type.cpp:176-184): normalised velocity (0, 0),
// acceleration = slideAccelerationLength * (0, 0) = (-0, -0), not rolling, no rolling time (System::append, type.cpp:510-512).
__device__ __forceinline__ BallRec load_rest_ball(const PipeWorkspace& w, int64_t g, const PhysConst& pc) {
    BallRec r;
    const float2 q0 = w.pos[g];
    const uint32_t m = w.meta[g];
    r.pos = mk(q0.x, q0.y); r.vel = mk(0.f, 0.f); r.vn = mk(0.f, 0.f); r.acc = rest_acc(pc);
    r.nextRolling = -1.0f;
    r.kind = meta_kind(m); r.rolling = false; r.lastRail = meta_last_rail(m); r.type = meta_type(m);
    return r;
}
__device__ __forceinline__ void store_ball(const PipeWorkspace& w, int64_t g, const BallRec& r) {
    w.pos[g] = make_float2(r.pos.x, r.pos.y);
    w.vel[g] = make_float2(r.vel.x, r.vel.y);
    w.vn[g] = make_float2(r.vn.x, r.vn.y);
    w.acc[g] = make_float2(r.acc.x, r.acc.y);
    w.nextRolling[g] = r.nextRolling;
    w.meta[g] = make_meta(r.kind, r.rolling, r.lastRail, r.type);
}
// System::append bookkeeping for one ball of the new layer (type.cpp:478-527); `cause` = rail index for K_RAIL
__device__ __forceinline__ void append_bookkeeping(BallRec& r, float deltaTime, int cause, const PhysConst& pc) {
    if (r.nextRolling >= 0.0f) { const float nr = r.nextRolling - deltaTime; r.nextRolling = nr > 0.0f ? nr : 0.0f; }
    if (r.kind == K_COLLISION) {
        r.lastRail = -1;
        r.nextRolling = (!r.rolling && !is_zero(r.vel)) ? time_to_rolling(r.vel, pc) : -1.0f;
    } else if (r.kind == K_REST || r.kind == K_POTTING) {
        r.nextRolling = -1.0f;
    } else if (r.kind == K_MOVING) {
        if (r.rolling) r.nextRolling = -1.0f;
    } else if (r.kind == K_RAIL) {
        r.lastRail = cause; r.nextRolling = -1.0f;
    }
}
__device__ __forceinline__ void offer_own(unsigned long long* key, const BallRec& r, int b) {
    offer(key, time_to_stop(r.acc, r.vel), make_code(b, CAT_REST, 0));
    if (r.nextRolling >= 0.0f) offer(key, r.nextRolling, make_code(b, CAT_ROLLING, 0));
}

__device__ __forceinline__ void apply_body(AppendShared& ash, const LaunchParams& p, const PipeWorkspace& w, int round, unsigned bid,
                                           unsigned nb) {
    const PhysConst pc = p_phys;
    const int BS = blockDim.x;
    const uint32_t nLive = w.counts[4 * round + 0], nDyn = w.counts[4 * round + 1];
    if (nLive == 0) return;
    const uint32_t* live = w.live + (size_t)(round & 1) * w.n;
    uint32_t* liveNext = w.live + (size_t)((round + 1) & 1) * w.n;
    const uint32_t* dyn = w.dyn + (size_t)(round & 1) * w.dynCap;
    uint32_t* dynNext = w.dyn + (size_t)((round + 1) & 1) * w.dynCap;
    const unsigned long long* keyCur = w.key + (size_t)(round % 3) * w.n;
    unsigned long long* keyNext = w.key + (size_t)((round + 1) % 3) * w.n;
    unsigned long long* keyStale = w.key + (size_t)((round + 2) % 3) * w.n;
    const float* Tcur = w.T + (size_t)(round & 1) * w.n;
    float* Tnext = w.T + (size_t)((round + 1) & 1) * w.n;
    uint32_t* liveCountNext = &w.counts[4 * (round + 1) + 0];
    uint32_t* dynCountNext = &w.counts[4 * (round + 1) + 1];
    const bool overflow = round + 1 > MAX_EVENTS;  // billiard_search.cpp:1376-1379 (every live strike has `round` events)
    // ---- the event of every live strike (createLayer :1957-1992) ----------------------------------------------------
    // One pass over both lists: tile i of the live strikes and tile i of the dynamic balls are handled by the same CTA
    // iteration (they touch different balls), so their memory round trips overlap and one list reservation serves both.
    const int64_t nMax = nLive > nDyn ? (int64_t)nLive : (int64_t)nDyn;
    const int64_t stride = (int64_t)nb * BS;
    for (int64_t base = (int64_t)bid * BS; base < nMax; base += stride) {
        const int64_t idx = base + threadIdx.x;
        bool appendLive = false, appendA = false, appendB = false;
        uint32_t entA = 0, entB = 0, sLive = 0;
        if (idx < nLive) {
            const int64_t s = live[idx];
            sLive = (uint32_t)s;
            const unsigned long long key = keyCur[s];
            keyStale[s] = ~0ull;
            const float T = Tcur[s];
            if (key == ~0ull) {
                mark_finished(p, s, F_NOEVENT, round, T);              // billiard_search.cpp:1464-1467
            } else if (overflow) {
                mark_finished(p, s, F_OVERFLOW, round, T);
            } else {
                const float dt = from_ordered_bits((uint32_t)(key >> 32));
                const uint32_t code = (uint32_t)key;
                const int evBall = (int)(code >> 16);
                const uint32_t cat = (code >> 8) & 0xffu;
                const int evPartner = (int)(code & 0xffu);
                const int evType = cat == CAT_REST ? E_REST : cat == CAT_ROLLING ? E_ROLLING : cat == CAT_RAIL ? E_RAIL_COLLISION
                                 : cat == CAT_POCKET ? E_POTTING : E_BALL_COLLISION;
                const float newT = T + dt;
                const float deltaTime = newT - T;  // type.cpp:475
                Tnext[s] = newT;
                uint32_t dynMask = w.dynMask[s], restMask = w.restMask[s];
                const int64_t gA = (int64_t)evBall * w.n + s;
                // the partner of a ball collision is loaded together with the event ball (one memory round trip, not two)
                const bool pairEv = cat == CAT_STATIC_PAIR || cat == CAT_DYNAMIC_PAIR;
                const int64_t gB = (int64_t)(pairEv ? evPartner : evBall) * w.n + s;
                BallRec A = load_ball(w, gA);
                BallRec B = cat == CAT_STATIC_PAIR ? load_rest_ball(w, gB, pc) : load_ball(w, gB);
                const v2 npos = advance_position(A.acc, A.vel, dt, A.pos);
                const v2 nvel = advance_velocity(A.acc, A.vel, dt);
                int cause = -1;
                A.pos = npos;
                if (evType == E_POTTING) {                 // :1716-1746
                    A.vel = mk(0.f, 0.f); A.rolling = false; A.kind = K_POTTING;
                } else if (evType == E_RAIL_COLLISION) {   // :1748-1786
                    const RailGeom& rg = w.table->rails[evPartner];
                    A.vel = rail_collision(nvel, mk(rg.nx, rg.ny), pc);
                    A.rolling = true; A.kind = K_RAIL; cause = evPartner;
                } else if (evType == E_REST) {             // :1788-1808
                    A.vel = mk(0.f, 0.f); A.rolling = false; A.kind = K_REST;
                } else if (evType == E_ROLLING) {          // :1810-1840
                    A.vel = nvel; A.rolling = true; A.kind = K_MOVING;
                } else {                                   // :1842-1905 createBallCollisionNodes
                    const v2 opos = advance_position(B.acc, B.vel, dt, B.pos);
                    const v2 ovel = advance_velocity(B.acc, B.vel, dt);
                    v2 o1, o2;
                    elastic_collision(npos, nvel, opos, ovel, A.rolling, B.rolling, pc, &o1, &o2);
                    A.vel = o1; A.kind = K_COLLISION;
                    B.pos = opos; B.vel = o2; B.kind = K_COLLISION;
                    derive_ball(B.vel, B.rolling, pc, &B.vn, &B.acc);
                    append_bookkeeping(B, deltaTime, -1, pc);
                    store_ball(w, gB, B);
                    offer_own(keyNext + s, B, evPartner);
                    dynMask |= 1u << evPartner; restMask &= ~(1u << evPartner);
                    appendB = true; entB = ((uint32_t)s << 5) | (uint32_t)evPartner;
                }
                derive_ball(A.vel, A.rolling, pc, &A.vn, &A.acc);
                append_bookkeeping(A, deltaTime, cause, pc);
                store_ball(w, gA, A);
                if (kind_dynamic(A.kind)) {
                    offer_own(keyNext + s, A, evBall);
                    appendA = true; entA = ((uint32_t)s << 5) | (uint32_t)evBall;
                } else {
                    dynMask &= ~(1u << evBall);
                    if (A.kind == K_REST) restMask |= 1u << evBall;
                }
                w.dynMask[s] = dynMask; w.restMask[s] = restMask;
                if (p.out.eventLog) {
                    int a = evBall, b = evPartner;
                    if (evType == E_BALL_COLLISION && a > b) { const int t = a; a = b; b = t; }
                    if (evType == E_REST || evType == E_ROLLING) b = -1;
                    p.out.eventLog[s * 32 + round] = make_int4(evType, a, b, __float_as_int(newT));
                }
                if (dynMask == 0) mark_finished(p, s, F_COMPLETED, round + 1, newT);
                else appendLive = true;
            }
        }
        // ---- every other dynamic ball: closed-form advance to the new layer (createNodes :1907-1955) -----------------
        bool keep = false;
        uint32_t ent = 0;
        if (idx < nDyn && !overflow) {
            ent = dyn[idx];
            const int64_t s = ent >> 5;
            const int b = (int)(ent & 31u);
            const unsigned long long key = keyCur[s];
            if (key != ~0ull) {
                const uint32_t code = (uint32_t)key;
                const int evBall = (int)(code >> 16);
                const uint32_t cat = (code >> 8) & 0xffu;
                const int evPartner = (int)(code & 0xffu);
                const bool pairEvent = cat == CAT_STATIC_PAIR || cat == CAT_DYNAMIC_PAIR;
                if (b != evBall && !(pairEvent && b == evPartner)) {
                    const float dt = from_ordered_bits((uint32_t)(key >> 32));
                    const float T = Tcur[s];
                    const float newT = T + dt;
                    const float deltaTime = newT - T;
                    const int64_t g = (int64_t)b * w.n + s;
                    BallRec r = load_ball(w, g);
                    const v2 nv = advance_velocity(r.acc, r.vel, dt);
                    r.pos = advance_position(r.acc, r.vel, dt, r.pos);
                    r.vel = nv; r.kind = K_MOVING;
                    derive_ball(r.vel, r.rolling, pc, &r.vn, &r.acc);
                    append_bookkeeping(r, deltaTime, -1, pc);
                    store_ball(w, g, r);
                    offer_own(keyNext + s, r, b);
                    keep = true;
                }
            }
        }
        {
            const bool pred[4] = {appendLive, appendA, appendB, keep};
            uint32_t* const counter[4] = {liveCountNext, dynCountNext, dynCountNext, dynCountNext};
            uint32_t slot[4];
            cta_reserve<4>(ash, pred, counter, slot);
            if (appendLive) liveNext[slot[0]] = sLive;
            if (appendA) dynNext[slot[1]] = entA;
            if (appendB) dynNext[slot[2]] = entB;
            if (keep) dynNext[slot[3]] = ent;
        }
    }
}

__global__ void __launch_bounds__(APPLY_BLOCK) apply_kernel(const LaunchParams* __restrict__ pp, PipeWorkspace w, int round) {
    if (w.pdlEarly) cudaTriggerProgrammaticLaunchCompletion();
    __shared__ AppendShared ash;
    cudaGridDependencySynchronize();
    apply_body(ash, *pp, w, round, blockIdx.x, gridDim.x);
}

#include "bcu_step.cuh"
#include "bcu_round.cuh"

// rules + simulationCost + result write-out (billiard_search.cpp:1388-1405, 1473-1532; billiard_snooker.cpp:1128-1153).
// One warp per strike, lane = ball.
__device__ __forceinline__ void finish_body(const LaunchParams& p, const PipeWorkspace& w, unsigned bid, unsigned nb) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int NB = p.states.nBalls;
    const int64_t warpsTotal = (int64_t)nb * (blockDim.x >> 5);
    unsigned long long events = 0;
    for (int64_t s = (int64_t)bid * (blockDim.x >> 5) + (threadIdx.x >> 5); s < p.nCand; s += warpsTotal) {
        const Candidate c = decode_candidate(p, s);
        uint32_t flags = p.out.flags[s];
        int kind = K_ABSENT, type = -1;
        float2 q = make_float2(0.f, 0.f);
        if (lane < NB) {
            // merged layout: the strike's last layer (layer nEvents) is in buffer nEvents & 1
            const StateBuf& sb = w.buf[w.merged ? (int)((flags >> 8) & 1u) : 0];
            const uint32_t m = sb.meta[(int64_t)lane * w.n + s];
            kind = meta_kind(m); type = meta_type(m);
            q = sb.pos[(int64_t)lane * w.n + s];
        }
        const unsigned pottedMask = __ballot_sync(FULL, kind == K_POTTING);
        const unsigned cuePotted = __ballot_sync(FULL, kind == K_POTTING && type == p.rules->cueType);
        const uint32_t expected = p.states.expectedMask[c.sIdx];
        const unsigned unexpected = __ballot_sync(FULL, kind == K_POTTING && !((expected >> (type & 31)) & 1u));
        uint64_t cost = ~0ull;
        if (flags & F_COMPLETED) {
            if (pottedMask != 0 && cuePotted == 0 && unexpected == 0) flags |= F_VALID;
            double pottingScore = 0.0;
            for (unsigned m = pottedMask; m; m &= m - 1) {   // ascending ball index, as the oracle sums
                const int b = __ffs(m) - 1;
                pottingScore += p.rules->score[__shfl_sync(FULL, type, b)];
            }
            const v2 v0 = mk(c.vx, c.vy);
            const float velocitySquared = dot(v0, v0);
            const double velocityCost = (double)(velocitySquared / (5000.0f * 5000.0f));
            const double pottingCost = 1.0 - (pottingScore < 1.0 ? pottingScore : 1.0);
            const double wp = 0.5 * pottingCost, wv = 0.5 * velocityCost;
            const double total = (wp + wv) < 1.0 ? (wp + wv) : 1.0;
            const double cap = 1.0 * (double)c.searchCost;
            cost = c.baseCost + (uint64_t)(total * cap);
        }
        if (lane == 0) {
            p.out.cost[s] = cost;
            p.out.flags[s] = flags;
            p.out.potted[s] = pottedMask;
            events += (flags >> 8) & 0xffu;
        }
        if (p.out.finalPos && lane < NB) reinterpret_cast<float2*>(p.out.finalPos)[s * NB + lane] = make_float2(q.x, q.y);
    }
    if (p.out.counters && lane == 0 && events) atomicAdd(&p.out.counters[0], events);
}
__global__ void __launch_bounds__(256) finish_kernel(const LaunchParams* __restrict__ pp, PipeWorkspace w) {
    if (w.pdlEarly) cudaTriggerProgrammaticLaunchCompletion();
    cudaGridDependencySynchronize();
    finish_body(*pp, w, blockIdx.x, gridDim.x);
}

// work counters from the per-round list sizes: [1] pair items [2] quartics solved [3] rail items [4] pocket items
// [5] pairs the reach bound skipped (or, in audit mode, marked)  [6] audit: marked jobs that would have mattered
__global__ void counters_kernel(const LaunchParams* __restrict__ pp, PipeWorkspace w) {
    const LaunchParams& p = *pp;
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned long long pairs = 0, quartics = 0, rails = 0, pockets = 0;
    for (int r = 0; r < ROUNDS; ++r) {
        const unsigned long long d = w.counts[4 * r + 1];
        if (w.counts[4 * r + 0] == 0) continue;
        pairs += d * p.states.nBalls; rails += d * p.table->nRails; pockets += d * p.table->nPockets;
        quartics += w.counts[4 * r + 2];
        quartics += w.counts[4 * r + 3];   // the queue's other end
    }
    atomicAdd(&p.out.counters[1], pairs); atomicAdd(&p.out.counters[2], quartics);
    atomicAdd(&p.out.counters[3], rails); atomicAdd(&p.out.counters[4], pockets);
    atomicAdd(&p.out.counters[6], (unsigned long long)w.counts[4 * ROUNDS + 2]);   // [5] (skipped pairs) is added by items_kernel
}

}  // namespace

void pipeline_upload_constants(cudaStream_t stream) {
    PhysConst pc = phys_const();
    cudaMemcpyToSymbolAsync(p_phys, &pc, sizeof(PhysConst), 0, cudaMemcpyHostToDevice, stream);
}

const uint32_t* pipeline_overflow_flag(const PipeWorkspace& w) { return w.counts + 4 * ROUNDS + 3; }

size_t pipeline_workspace_bytes(int64_t n, int nBalls, int jobsPerStrike, PipeWorkspace* layout) {
    auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
    PipeWorkspace w{};
    w.n = n;
    w.dynCap = (size_t)n * nBalls;
    w.jobCap = (size_t)n * jobsPerStrike;
    if (w.jobCap < 4096) w.jobCap = 4096;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off += align(bytes); return o; };
    const size_t nb = (size_t)n * nBalls;
    const size_t oPos = take(nb * sizeof(float2)), oVel = take(nb * sizeof(float2)), oVn = take(nb * sizeof(float2));
    const size_t oAcc = take(nb * sizeof(float2)), oNr = take(nb * sizeof(float));
    const size_t oMeta = take(nb * sizeof(uint32_t));
    const size_t oLc = take(nb * sizeof(uint32_t));
    const size_t oKey = take((size_t)3 * n * sizeof(unsigned long long));
    const size_t oT = take((size_t)2 * n * sizeof(float));
    const size_t oDm = take((size_t)n * sizeof(uint32_t)), oRm = take((size_t)n * sizeof(uint32_t));
    const size_t oPos2 = take(nb * sizeof(float2)), oVel2 = take(nb * sizeof(float2)), oVn2 = take(nb * sizeof(float2));
    const size_t oAcc2 = take(nb * sizeof(float2)), oNr2 = take(nb * sizeof(float)), oMeta2 = take(nb * sizeof(uint32_t));
    const size_t oDm2 = take((size_t)n * sizeof(uint32_t)), oRm2 = take((size_t)n * sizeof(uint32_t));
    const size_t oStop = take((size_t)2 * n * sizeof(uint32_t));
    const size_t oLive = take((size_t)2 * n * sizeof(uint32_t));
    const size_t oDyn = take((size_t)2 * w.dynCap * sizeof(uint32_t));
    const size_t oCounts = take((size_t)4 * (ROUNDS + 1) * sizeof(uint32_t));
    const size_t oCo = take((size_t)5 * w.jobCap * sizeof(float)), oCode = take(w.jobCap * sizeof(uint32_t));
    if (layout) {
        unsigned char* b = reinterpret_cast<unsigned char*>(layout->base);
        w.base = layout->base;
        w.pos = reinterpret_cast<float2*>(b + oPos); w.vel = reinterpret_cast<float2*>(b + oVel);
        w.vn = reinterpret_cast<float2*>(b + oVn); w.acc = reinterpret_cast<float2*>(b + oAcc);
        w.nextRolling = reinterpret_cast<float*>(b + oNr); w.meta = reinterpret_cast<uint32_t*>(b + oMeta); w.lineCand = reinterpret_cast<uint32_t*>(b + oLc);
        w.key = reinterpret_cast<unsigned long long*>(b + oKey); w.T = reinterpret_cast<float*>(b + oT);
        w.dynMask = reinterpret_cast<uint32_t*>(b + oDm); w.restMask = reinterpret_cast<uint32_t*>(b + oRm);
        w.live = reinterpret_cast<uint32_t*>(b + oLive); w.dyn = reinterpret_cast<uint32_t*>(b + oDyn);
        w.counts = reinterpret_cast<uint32_t*>(b + oCounts);
        w.jobCo = reinterpret_cast<float*>(b + oCo); w.jobCode = reinterpret_cast<uint32_t*>(b + oCode);
        w.buf[0] = StateBuf{w.pos, w.vel, w.vn, w.acc, w.nextRolling, w.meta};
        w.buf[1] = StateBuf{reinterpret_cast<float2*>(b + oPos2), reinterpret_cast<float2*>(b + oVel2), reinterpret_cast<float2*>(b + oVn2),
                            reinterpret_cast<float2*>(b + oAcc2), reinterpret_cast<float*>(b + oNr2), reinterpret_cast<uint32_t*>(b + oMeta2)};
        w.maskDyn[0] = w.dynMask; w.maskRest[0] = w.restMask;
        w.maskDyn[1] = reinterpret_cast<uint32_t*>(b + oDm2); w.maskRest[1] = reinterpret_cast<uint32_t*>(b + oRm2);
        w.stopped[0] = reinterpret_cast<uint32_t*>(b + oStop); w.stopped[1] = w.stopped[0] + n;
        *layout = w;
    }
    return off;
}

int pipeline_launch_count(const PipeLaunchCfg& cfg) { return cfg.layout != 0 ? 2 + (ROUNDS + 1) + ROUNDS : 2 + 3 * ROUNDS; }

// One launch of the batch path over p->nCand <= w.n strikes.  `p` is a DEVICE pointer (the kernels read their parameters
// from memory, so a captured CUDA graph of this sequence stays valid from call to call); all grids are persistent and
// sized by the workspace capacity, so nothing in the launch sequence depends on the candidate count.
cudaError_t launch_pipeline(const LaunchParams* p, int nBalls, int nRails, int nPockets, bool withCounters,
                            const PipeWorkspace& wsIn, int sms, cudaStream_t stream, bool resetOverflow, PipeLaunchCfg& cfg) {
    PipeWorkspace w = wsIn;
    w.fused = cfg.layout == 1;
    w.merged = cfg.layout == 2;
    cudaError_t e;
    if ((e = cudaMemsetAsync(w.key, 0xff, (size_t)3 * w.n * sizeof(unsigned long long), stream)) != cudaSuccess) return e;
    // the overflow flag is the last word of counts[]: sticky across the parts that share this workspace
    const size_t countWords = (size_t)4 * (ROUNDS + 1) - (resetOverflow ? 0 : 1);
    if ((e = cudaMemsetAsync(w.counts, 0, countWords * sizeof(uint32_t), stream)) != cudaSuccess) return e;
    const size_t itemSmem = (size_t)ITEM_BLOCK * (2 * nBalls) * sizeof(uint16_t);
    (void)nRails; (void)nPockets;
    const int64_t initBlocks = (w.n + 255) / 256, finBlocks = (w.n + 7) / 8;
    // BCU_ROUND_TIMING=1 (diagnosis only): CUDA events around every kernel, per-round times on stderr
    const bool timing = cfg.roundTiming;
    std::vector<cudaEvent_t> ev;
    auto mark = [&]() {
        if (!timing) return;
        cudaEvent_t x; cudaEventCreate(&x); cudaEventRecord(x, stream); ev.push_back(x);
    };
    // persistent grids: exactly one resident wave of every kernel
    int& occItems = cfg.occItems; int& occQuartic = cfg.occQuartic; int& occApply = cfg.occApply;
    size_t& occSmem = cfg.occSmem;
    const size_t roundSmem = (size_t)2 * RK_SLOTS * nBalls * sizeof(uint16_t);
    const int lanes = cfg.stepLanes == 4 ? 4 : cfg.stepLanes == 16 ? 16 : 8;
    const size_t stepSmem = (size_t)(STEP_BLOCK / lanes) * step_group_bytes(nBalls);
    auto stepKernel = lanes == 4 ? step_kernel<4> : lanes == 16 ? step_kernel<16> : step_kernel<8>;
    if (occSmem != itemSmem) {
        if (stepSmem > 48 * 1024) cudaFuncSetAttribute(stepKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stepSmem);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cfg.occStep, stepKernel, STEP_BLOCK, stepSmem);
        if (cfg.occStep < 1) cfg.occStep = 1;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cfg.occRound, round_kernel, RK_BLOCK, roundSmem);
        if (cfg.occRound < 1) cfg.occRound = 1;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occItems, items_kernel, ITEM_BLOCK, itemSmem);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occQuartic, quartic_kernel, QUARTIC_BLOCK, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occApply, apply_kernel, APPLY_BLOCK, 0);
        if (occItems < 1) occItems = 1;
        if (occQuartic < 1) occQuartic = 1;
        if (occApply < 1) occApply = 1;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cfg.occItemsDense, items_kernel_dense, ITEM_BLOCK, itemSmem);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&cfg.occQuarticWide, quartic_kernel_wide, QUARTIC_WIDE_BLOCK, 0);
        if (cfg.occItemsDense < 1) cfg.occItemsDense = 1;
        if (cfg.occQuarticWide < 1) cfg.occQuarticWide = 1;
        occSmem = itemSmem;
    }
    // Programmatic dependent launch: every kernel after init is launched with the stream-serialisation attribute and signals
    // cudaTriggerProgrammaticLaunchCompletion() at once, so the CTAs of kernel k+1 are placed (and run their prologue: table
    // geometry to shared memory) while kernel k drains; they read nothing kernel k writes before cudaGridDependencySynchronize().
    // A round is three latency-bound kernels, so launch gaps are a visible share of a level.  BCU_NO_PDL=1 switches it off.
    const int pdlMode = cfg.pdlMode;
    const bool pdl = pdlMode > 0;
    w.pdlEarly = pdlMode > 1;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    auto cfgOf = [&](unsigned grid, unsigned block, size_t smem) {
        cudaLaunchConfig_t c{};
        grid = (unsigned)((uint64_t)grid * (uint64_t)cfg.gridPct / 100u);
        if (grid < 1) grid = 1;
        c.gridDim = dim3(grid); c.blockDim = dim3(block); c.dynamicSmemBytes = smem; c.stream = stream;
        c.attrs = attr; c.numAttrs = pdl ? 1 : 0;
        return c;
    };
    mark();
    init_kernel<<<(unsigned)(initBlocks < sms * 8 ? initBlocks : sms * 8), 256, 0, stream>>>(p, w);
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    mark();
    if (w.merged) {
        // a round is two kernels: round k applies event k of every live strike (reading layer k - 1, writing layer k) and finds
        // the candidates of layer k in the same thread-per-dynamic-ball pass; quartic k solves the pair jobs
        for (int k = 0; k <= ROUNDS; ++k) {
            cudaLaunchConfig_t cr = cfgOf(sms * cfg.occRound, RK_BLOCK, roundSmem);
            if ((e = cudaLaunchKernelEx(&cr, round_kernel, p, w, k)) != cudaSuccess) return e;
            mark();
            if (k == ROUNDS) break;
            cudaLaunchConfig_t cq = cfgOf(sms * occQuartic, QUARTIC_BLOCK, 0);
            if ((e = cudaLaunchKernelEx(&cq, quartic_kernel, w, k)) != cudaSuccess) return e;
            mark();
        }
    }
    if (w.fused) {
        // a round is two kernels: step k applies event k of every live strike and collects the candidates of the new layer
        // (rest / rolling / rail / pocket times -> the strike's key, ball pairs -> quartic jobs); quartic k solves the jobs
        for (int k = 0; k <= ROUNDS; ++k) {
            cudaLaunchConfig_t cs = cfgOf(sms * cfg.occStep, STEP_BLOCK, stepSmem);
            if ((e = cudaLaunchKernelEx(&cs, stepKernel, p, w, k)) != cudaSuccess) return e;
            mark();
            if (k == ROUNDS) break;
            cudaLaunchConfig_t cq = cfgOf(sms * occQuartic, QUARTIC_BLOCK, 0);
            if ((e = cudaLaunchKernelEx(&cq, quartic_kernel, w, k)) != cudaSuccess) return e;
            mark();
        }
    }
    // Parts of 65 536 strikes and more (cfg.denseMinItems / denseMinQuartic) take the kernels for large parts: items_kernel_dense
    // (the same body compiled for more resident CTAs per SM) and quartic_kernel_wide (CTAs whose warps walk the solver side by
    // side).  Measured (profiles/r02_dense_sweep.txt, profiles/r02_ab_experiments.txt): 1000 tables x 4096 strikes (parts of 2^20)
    // 116.6 M strikes/s against 93.7 M with the small-part kernels; parts of 65 536: +2-4 %; parts of 32 768 (the 65 536-strike
    // level, latency-bound rounds) lose 1-5 % with them.  Same code, same results.
    // With the reach bound off a round has about twice the quartic jobs, and the wide kernel already pays on parts of 32 768
    // strikes (the benched level: 31.2 M against 30.0 M strikes/s; with the bound on 40.5 against 40.7 M).
    const bool denseI = w.n >= cfg.denseMinItems, denseQ = w.n >= (w.pruneMode == 1 ? cfg.denseMinQuartic : cfg.denseMinQuartic / 2);
    for (int r = 0; r < ROUNDS && !w.fused && !w.merged; ++r) {
        cudaLaunchConfig_t ci = cfgOf(sms * (denseI ? cfg.occItemsDense : occItems), ITEM_BLOCK, itemSmem);
        if ((e = cudaLaunchKernelEx(&ci, denseI ? items_kernel_dense : items_kernel, p, w, r)) != cudaSuccess) return e;
        mark();
        cudaLaunchConfig_t cq = denseQ ? cfgOf(sms * cfg.occQuarticWide, QUARTIC_WIDE_BLOCK, 0) : cfgOf(sms * occQuartic, QUARTIC_BLOCK, 0);
        if ((e = cudaLaunchKernelEx(&cq, denseQ ? quartic_kernel_wide : quartic_kernel, w, r)) != cudaSuccess) return e;
        mark();
        if (timing && cfg.quarticTwice) {   // diagnosis: the same jobs again (offers are idempotent) with the solver's code still cached
            if ((e = cudaLaunchKernelEx(&cq, denseQ ? quartic_kernel_wide : quartic_kernel, w, r)) != cudaSuccess) return e;
            mark();
        }
        cudaLaunchConfig_t ca = cfgOf(sms * occApply, APPLY_BLOCK, 0);
        if ((e = cudaLaunchKernelEx(&ca, apply_kernel, p, w, r)) != cudaSuccess) return e;
        mark();
    }
    {
        cudaLaunchConfig_t cf = cfgOf((unsigned)(finBlocks < sms * 8 ? finBlocks : sms * 8), 256, 0);
        if ((e = cudaLaunchKernelEx(&cf, finish_kernel, p, w)) != cudaSuccess) return e;
    }
    mark();
    if (timing && (w.fused || w.merged)) {
        cudaStreamSynchronize(stream);
        std::vector<uint32_t> counts(4 * (ROUNDS + 1));
        cudaMemcpy(counts.data(), w.counts, counts.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost);
        auto ms = [&](size_t i) { float t = 0; cudaEventElapsedTime(&t, ev[i], ev[i + 1]); return t * 1000.0f; };
        float sum[2] = {0, 0};
        fprintf(stderr, "[bcu timing] init %.1f us\n", ms(0));
        for (int k = 0; k <= ROUNDS; ++k) {
            const float ts = ms(1 + 2 * k), tq = k < ROUNDS ? ms(2 + 2 * k) : 0.0f;
            fprintf(stderr, "[bcu timing] step %2d live %7u jobs %7u : step %6.1f quartic %6.1f us\n", k, counts[4 * k],
                    k < ROUNDS ? counts[4 * k + 2] : 0u, ts, tq);
            sum[0] += ts; sum[1] += tq;
        }
        float tot = 0; cudaEventElapsedTime(&tot, ev.front(), ev.back());
        fprintf(stderr, "[bcu timing] finish %.1f us; step %.0f quartic %.0f; total %.0f us\n", ms(ev.size() - 2), sum[0], sum[1], tot * 1000.0f);
        for (auto x : ev) cudaEventDestroy(x);
    } else if (timing) {
        cudaStreamSynchronize(stream);
        std::vector<uint32_t> counts(4 * (ROUNDS + 1));
        cudaMemcpy(counts.data(), w.counts, counts.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost);
        auto ms = [&](size_t i) { float t = 0; cudaEventElapsedTime(&t, ev[i], ev[i + 1]); return t * 1000.0f; };
        float sum[3] = {0, 0, 0}, again = 0;
        fprintf(stderr, "[bcu timing] init %.1f us\n", ms(0));
        const int per = cfg.quarticTwice ? 4 : 3;   // marks per round
        for (int r = 0; r < ROUNDS; ++r) {
            const float ti = ms(1 + per * r), tq = ms(2 + per * r), ta = ms(per + per * r);
            fprintf(stderr, "[bcu timing] round %2d live %7u dyn %7u jobs %7u : items %6.1f quartic %6.1f apply %6.1f us", r,
                    counts[4 * r], counts[4 * r + 1], counts[4 * r + 2] + counts[4 * r + 3], ti, tq, ta);
            if (cfg.quarticTwice) { fprintf(stderr, "  quartic again %6.1f", ms(3 + per * r)); again += ms(3 + per * r); }
            fprintf(stderr, "\n");
            sum[0] += ti; sum[1] += tq; sum[2] += ta;
        }
        if (cfg.quarticTwice) fprintf(stderr, "[bcu timing] quartic launched twice per round: first %.0f us, again (code cached) %.0f us\n", sum[1], again);
        float tot = 0; cudaEventElapsedTime(&tot, ev.front(), ev.back());
        fprintf(stderr, "[bcu timing] finish %.1f us; items %.0f quartic %.0f apply %.0f; total %.0f us\n", ms(ev.size() - 2), sum[0],
                sum[1], sum[2], tot * 1000.0f);
        for (auto x : ev) cudaEventDestroy(x);
    }
    if (withCounters) counters_kernel<<<1, 32, 0, stream>>>(p, w);
    return cudaGetLastError();
}

}  // namespace bcu
