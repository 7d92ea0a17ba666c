// K1 v4: the persistent tile kernel (the batch path of bcu_evaluate / bcu_evaluate_grid).
//
// Timing the event-synchronous pipeline (bcu_pipeline.cu) round by round showed that a level is bound by the CHAIN, not
// by throughput: 31 rounds x three dependent kernels of 15-30 us each is ~1.9 ms although the arithmetic of all 65 536
// strikes overlaps into ~0.4 ms.  This kernel keeps the pipeline's phases and its dense, stream-compacted mapping but
// replaces every kernel boundary by a __syncthreads():
//
//   * a CTA owns S strike slots; their table states live in SHARED memory, ball-major ([ball][slot]);
//   * per round the CTA (A) lists the dynamic balls of its live slots, (B) runs the cheap first-stage rail / pocket /
//     partner tests with K threads per dynamic ball and compacts the survivors into shared-memory queues, (C) runs the
//     second stages -- rail / pocket quadratics, moving-vs-moving pre-checks, and one thread per surviving ball pair for
//     the quantised coefficients + FP64 closed-form quartic -- on dense threads, (D) applies the winning event of every
//     slot and advances the other dynamic balls in closed form;
//   * times go to the slot's 64-bit (ordered time bits << 32 | evaluation-order code) key with a shared-memory
//     atomicMin: the minimum is the reference's strict-'<' first-evaluated-wins choice (billiard_search.cpp:2002-2013);
//   * a slot whose strike reached its final layer is costed, written out and refilled from a global candidate counter,
//     so the CTAs of an SM sit in different phases and hide each other's latencies, and long strikes (up to MAX_EVENTS
//     rounds) do not hold short ones back.
//
// Reference path: billiard_search.cpp:1358-1471 (simulate), :1550-1573 (nextEvent), :1957-1992 (createLayer),
// :1473-1532 (simulationCost), type.cpp:467-531 (System::append); billiard_physics.cpp throughout.
// Compiled with -fmad=false: all arithmetic is un-contracted IEEE.
#include <cuda_runtime.h>

#include "bcu_kernels.cuh"
#include "bcu_physics.cuh"

namespace bcu {

namespace {

__constant__ TableGeom t_table;
__constant__ RulesData t_rules;
__constant__ PhysConst t_phys;

constexpr int TILE_THREADS = 256;
constexpr int TILE_K = 4;                           // threads per dynamic ball in the first stage
constexpr int TILE_CHUNK = TILE_THREADS / TILE_K;   // dynamic balls per first-stage pass
constexpr int TILE_PAIR_CAP = 4096;                 // ball-pair queue (flushed through the quartic when it could overflow)

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

// Appends the set bits of `mask` (as base | bit index) to a shared-memory queue; all 32 lanes must call.
__device__ __forceinline__ void push_mask(uint32_t mask, uint32_t base, uint16_t* queue, int* counter) {
    const int lane = threadIdx.x & 31;
    const int cnt = __popc(mask);
    int incl = cnt;
    #pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += v;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    if (total == 0) return;
    int at = 0;
    if (lane == 31) at = atomicAdd(counter, total);
    at = __shfl_sync(0xffffffffu, at, 31) + incl - cnt;
    while (mask) {
        const int bit = __ffs(mask) - 1;
        mask &= mask - 1;
        queue[at++] = (uint16_t)(base | (uint32_t)bit);
    }
}

// Shared-memory view of a CTA's slots.  Ball arrays are [ball][slot]; acceleration is not stored (it is
// accelerationLength * velocityNormed, type.cpp:176-184, recomputed where needed from the rolling flag).
struct Tile {
    float2* pos;
    float2* vel;
    float2* vn;
    float* nextRolling;
    uint32_t* meta;
    unsigned long long* key;   // [2][S]
    float* T;
    uint32_t* dynMask;
    uint32_t* restMask;
    int32_t* cand;             // candidate index of the launch, -1: slot empty
    int32_t* nEvents;
    uint32_t* fin;             // 0: running, else the end flags
    uint16_t* entries;         // [S * NB] slot << 5 | ball, the dynamic balls of the round
    uint16_t* qRails;          // [CHUNK * R]    chunk entry << 5 | rail
    uint16_t* qPockets;        // [CHUNK * P]
    uint16_t* qDyn;            // [CHUNK * NB]   chunk entry << 5 | partner
    uint16_t* qPairs;          // [TILE_PAIR_CAP] round entry << 5 | partner
    int S, NB;
};

struct TileCounts {
    int entries, rails, pockets, dynPairs, pairs;
    int live, exhausted;
};

__host__ __device__ inline size_t tile_smem_bytes(int S, int NB, int R, int P) {
    size_t b = 0;
    b += (size_t)NB * S * (3 * sizeof(float2) + sizeof(float) + sizeof(uint32_t));
    b += (size_t)S * (2 * sizeof(unsigned long long) + sizeof(float) + 5 * sizeof(uint32_t));
    b += (size_t)S * NB * sizeof(uint16_t);
    b += (size_t)TILE_CHUNK * (R + P + NB) * sizeof(uint16_t);
    b += (size_t)TILE_PAIR_CAP * sizeof(uint16_t);
    return (b + 15) & ~(size_t)15;
}

__device__ __forceinline__ Tile carve_tile(unsigned char* raw, int S, int NB, int R, int P) {
    Tile t;
    t.S = S; t.NB = NB;
    unsigned char* q = raw;
    t.key = reinterpret_cast<unsigned long long*>(q); q += (size_t)2 * S * sizeof(unsigned long long);
    t.pos = reinterpret_cast<float2*>(q); q += (size_t)NB * S * sizeof(float2);
    t.vel = reinterpret_cast<float2*>(q); q += (size_t)NB * S * sizeof(float2);
    t.vn = reinterpret_cast<float2*>(q); q += (size_t)NB * S * sizeof(float2);
    t.nextRolling = reinterpret_cast<float*>(q); q += (size_t)NB * S * sizeof(float);
    t.meta = reinterpret_cast<uint32_t*>(q); q += (size_t)NB * S * sizeof(uint32_t);
    t.T = reinterpret_cast<float*>(q); q += (size_t)S * sizeof(float);
    t.dynMask = reinterpret_cast<uint32_t*>(q); q += (size_t)S * sizeof(uint32_t);
    t.restMask = reinterpret_cast<uint32_t*>(q); q += (size_t)S * sizeof(uint32_t);
    t.cand = reinterpret_cast<int32_t*>(q); q += (size_t)S * sizeof(int32_t);
    t.nEvents = reinterpret_cast<int32_t*>(q); q += (size_t)S * sizeof(int32_t);
    t.fin = reinterpret_cast<uint32_t*>(q); q += (size_t)S * sizeof(uint32_t);
    t.entries = reinterpret_cast<uint16_t*>(q); q += (size_t)S * NB * sizeof(uint16_t);
    t.qRails = reinterpret_cast<uint16_t*>(q); q += (size_t)TILE_CHUNK * R * sizeof(uint16_t);
    t.qPockets = reinterpret_cast<uint16_t*>(q); q += (size_t)TILE_CHUNK * P * sizeof(uint16_t);
    t.qDyn = reinterpret_cast<uint16_t*>(q); q += (size_t)TILE_CHUNK * NB * sizeof(uint16_t);
    t.qPairs = reinterpret_cast<uint16_t*>(q);
    return t;
}

__device__ __forceinline__ v2 ball_acc(v2 vn, bool rolling, const PhysConst& pc) {
    return (rolling ? pc.accLen : pc.slideAccLen) * vn;   // accelerationByNormed (billiard_physics.cpp:271-273)
}

// ---- slot refill: toInputLayer (billiard_search.cpp:760-795) + System::append of layer 0; one thread per slot ---------
__device__ __forceinline__ void refill_slot(const LaunchParams& p, const Tile& t, int slot, int64_t cand, unsigned long long* key,
                                            const PhysConst& pc) {
    const int S = t.S, NB = t.NB;
    const Candidate c = decode_candidate(p, cand);
    uint32_t dynMask = 0, restMask = 0;
    unsigned long long best = ~0ull;
    for (int b = 0; b < NB; ++b) {
        const int type = p.states.type[(int64_t)c.sIdx * NB + b];
        int kind = K_ABSENT;
        v2 pos = mk(0.f, 0.f), vel = mk(0.f, 0.f), vn, acc;
        float nextRolling = -1.0f;
        if (type >= 0) {
            const float2 q = reinterpret_cast<const float2*>(p.states.pos)[(int64_t)c.sIdx * NB + b];
            pos = mk(q.x, q.y);
            if (type == t_rules.cueType) {
                vel = mk(c.vx, c.vy);
                kind = K_SHOT;
                nextRolling = time_to_rolling(vel, pc);  // type.cpp:502-506
                dynMask |= 1u << b;
            } else {
                kind = K_REST;
                restMask |= 1u << b;
            }
        }
        derive_ball(vel, false, pc, &vn, &acc);
        const int g = b * S + slot;
        t.pos[g] = make_float2(pos.x, pos.y);
        t.vel[g] = make_float2(vel.x, vel.y);
        t.vn[g] = make_float2(vn.x, vn.y);
        t.nextRolling[g] = nextRolling;
        t.meta[g] = make_meta(kind, false, -1, type);
        if (kind == K_SHOT) {   // this thread owns the slot's key: plain minimum
            const float ts = time_to_stop(acc, vel);
            if (ts == ts) {
                const unsigned long long k = ((unsigned long long)ordered_bits(ts) << 32) | make_code(b, CAT_REST, 0);
                best = k < best ? k : best;
            }
            if (nextRolling >= 0.0f) {
                const unsigned long long k = ((unsigned long long)ordered_bits(nextRolling) << 32) | make_code(b, CAT_ROLLING, 0);
                best = k < best ? k : best;
            }
        }
    }
    *key = best;
    t.dynMask[slot] = dynMask; t.restMask[slot] = restMask; t.T[slot] = 0.0f;
    t.cand[slot] = (int32_t)cand; t.nEvents[slot] = 0;
    t.fin[slot] = dynMask == 0 ? F_COMPLETED : 0u;   // nothing moves: final at once
}

// ---- rules + simulationCost + result write-out of one finished slot (billiard_search.cpp:1388-1405, 1473-1532;
//      billiard_snooker.cpp:1128-1153); one thread per slot ----------------------------------------------------------------
__device__ __forceinline__ void finish_slot(const LaunchParams& p, const Tile& t, int slot) {
    const int S = t.S, NB = t.NB;
    const int64_t cand = t.cand[slot];
    uint32_t flags = t.fin[slot];
    const Candidate c = decode_candidate(p, cand);
    const uint32_t expected = p.states.expectedMask[c.sIdx];
    uint32_t pottedMask = 0;
    bool cuePotted = false, unexpected = false;
    double pottingScore = 0.0;
    for (int b = 0; b < NB; ++b) {   // ascending ball index, as the oracle sums
        const uint32_t m = t.meta[b * S + slot];
        if (meta_kind(m) == K_POTTING) {
            const int type = meta_type(m);
            pottedMask |= 1u << b;
            cuePotted |= type == t_rules.cueType;
            unexpected |= !((expected >> (type & 31)) & 1u);
            pottingScore += t_rules.score[type];
        }
    }
    uint64_t cost = ~0ull;
    if (flags & F_COMPLETED) {
        if (pottedMask != 0 && !cuePotted && !unexpected) flags |= F_VALID;
        const v2 v0 = mk(c.vx, c.vy);
        const float velocitySquared = dot(v0, v0);
        const double velocityCost = (double)(velocitySquared / (5000.0f * 5000.0f));
        const double pottingCost = 1.0 - (pottingScore < 1.0 ? pottingScore : 1.0);
        const double wp = 0.5 * pottingCost, wv = 0.5 * velocityCost;
        const double total = (wp + wv) < 1.0 ? (wp + wv) : 1.0;
        const double cap = 1.0 * (double)c.searchCost;
        cost = c.baseCost + (uint64_t)(total * cap);
    }
    p.out.cost[cand] = cost;
    p.out.flags[cand] = flags | ((uint32_t)t.nEvents[slot] << 8);
    p.out.potted[cand] = pottedMask;
    p.out.finalTime[cand] = t.T[slot];
    if (p.out.finalPos) {
        float2* dst = reinterpret_cast<float2*>(p.out.finalPos) + cand * NB;
        for (int b = 0; b < NB; ++b) dst[b] = t.pos[b * S + slot];
    }
}

// ---- second stage of one surviving ball pair: quantised coefficients (billiard_physics.cpp:326-337) + quartic ----------
__device__ __forceinline__ void solve_pair(const Tile& t, unsigned long long* keyCur, uint32_t code, float diameter,
                                           const PhysConst& pc) {
    const int S = t.S;
    const uint32_t ent = t.entries[code >> 5];
    const int o = (int)(code & 31u), b = (int)(ent & 31u), slot = (int)(ent >> 5);
    const int g1 = b * S + slot, g2 = o * S + slot;
    const float2 x1 = t.pos[g1], y1 = t.vel[g1], n1 = t.vn[g1], x2 = t.pos[g2], y2 = t.vel[g2], n2 = t.vn[g2];
    const v2 a1 = ball_acc(mk(n1.x, n1.y), meta_rolling(t.meta[g1]), pc);
    const v2 a2 = ball_acc(mk(n2.x, n2.y), meta_rolling(t.meta[g2]), pc);
    float co[5];
    collision_coefficients(a1, mk(y1.x, y1.y), mk(x1.x, x1.y), a2, mk(y2.x, y2.y), mk(x2.x, x2.y), diameter, co);
    float tc;
    if (collision_time_from_coefficients(co, &tc)) {
        const bool oStatic = (t.restMask[slot] >> o) & 1u;
        offer(keyCur + slot, tc, make_code(b, oStatic ? CAT_STATIC_PAIR : CAT_DYNAMIC_PAIR, o));
    }
}

__global__ void __launch_bounds__(TILE_THREADS, 3) tile_kernel(const LaunchParams* __restrict__ pp, unsigned long long* nextCand,
                                                               int S) {
    extern __shared__ __align__(16) unsigned char smemRaw[];
    __shared__ TileCounts cnt;
    __shared__ RailGeom sRails[MAX_RAILS];
    __shared__ float sDetQV[MAX_RAILS];
    __shared__ PocketGeom sPockets[MAX_POCKETS];
    const LaunchParams& p = *pp;
    const unsigned FULL = 0xffffffffu;
    const int tid = threadIdx.x, lane = tid & 31;
    const int R = t_table.nRails, P = t_table.nPockets, NB = p.states.nBalls;
    const Tile t = carve_tile(smemRaw, S, NB, R, P);
    const PhysConst pc = t_phys;
    const float diameter = t_table.diameter;
    const float dd = diameter * diameter;
    const int64_t nCand = p.nCand;

    for (int i = tid; i < R; i += TILE_THREADS) {
        const RailGeom rg = t_table.rails[i];
        sRails[i] = rg;
        sDetQV[i] = rg.sx * rg.dy - rg.sy * rg.dx;  // detQV of lineIntersectsLine (:651-723): depends on the rail only
    }
    for (int i = tid; i < P; i += TILE_THREADS) sPockets[i] = t_table.pockets[i];
    for (int i = tid; i < S; i += TILE_THREADS) { t.cand[i] = -1; t.fin[i] = 0; t.key[i] = ~0ull; t.key[S + i] = ~0ull; }
    if (tid == 0) { cnt.exhausted = 0; cnt.pairs = 0; }
    int cur = 0;   // key buffer of the current round
    unsigned long long cEvents = 0, cPairs = 0, cQuartics = 0, cRails = 0, cPockets = 0;   // thread 0 only
    __syncthreads();

    for (;;) {
        unsigned long long* keyCur = t.key + cur * S;
        unsigned long long* keyNext = t.key + (cur ^ 1) * S;

        // ---- (0) refill empty slots, finish what is final at once, count live slots ------------------------------------
        if (tid == 0) { cnt.live = 0; cnt.entries = 0; }
        __syncthreads();
        if (tid < ((S + 31) & ~31)) {   // whole warps
            const bool empty = tid < S && t.cand[tid] < 0 && !cnt.exhausted;
            const unsigned m = __ballot_sync(FULL, empty);
            if (m) {
                unsigned long long first = 0;
                if (lane == __ffs(m) - 1) first = atomicAdd(nextCand, (unsigned long long)__popc(m));
                first = __shfl_sync(FULL, first, __ffs(m) - 1);
                if (empty) {
                    const long long c = (long long)(first + __popc(m & ((1u << lane) - 1u)));
                    if (c < nCand) refill_slot(p, t, tid, c, keyCur + tid, pc);
                    else cnt.exhausted = 1;
                }
            }
        }
        __syncthreads();
        if (tid < S && t.cand[tid] >= 0) {
            if (t.fin[tid] != 0) {     // a refilled strike without a dynamic ball
                finish_slot(p, t, tid);
                t.cand[tid] = -1; t.fin[tid] = 0; keyCur[tid] = ~0ull;
                atomicAdd(&cnt.live, 1);   // the slot is refilled on the next pass: keep the loop going
            } else {
                atomicAdd(&cnt.live, 1);
            }
        }
        __syncthreads();
        if (cnt.live == 0) break;

        // ---- (A) the dynamic balls of the round --------------------------------------------------------------------------
        if (tid < ((S + 31) & ~31)) {
            const uint32_t m = (tid < S && t.cand[tid] >= 0) ? t.dynMask[tid] : 0u;
            // reuse push_mask: entries[] takes slot << 5 | ball
            push_mask(m, (uint32_t)tid << 5, t.entries, &cnt.entries);
        }
        __syncthreads();
        const int nEnt = cnt.entries;

        // ---- (B) + (C) first stages in chunks of TILE_CHUNK dynamic balls, second stages on the compacted survivors ----
        for (int c0 = 0; c0 < nEnt; c0 += TILE_CHUNK) {
            if (tid == 0) { cnt.rails = 0; cnt.pockets = 0; cnt.dynPairs = 0; }
            // the pair queue must be able to take this chunk's worst case; otherwise solve what is queued first
            if (cnt.pairs + TILE_CHUNK * NB > TILE_PAIR_CAP) {   // block-uniform (cnt.pairs is stable here)
                const int np = cnt.pairs;
                for (int i = tid; i < np; i += TILE_THREADS) solve_pair(t, keyCur, t.qPairs[i], diameter, pc);
                if (tid == 0) cQuartics += np;
                __syncthreads();
                if (tid == 0) cnt.pairs = 0;
            }
            __syncthreads();

            // stage 1: K threads per ball, thread `sub` takes every K-th rail / pocket / partner
            {
                const int ce = tid / TILE_K, sub = tid % TILE_K, e = c0 + ce;
                uint32_t railMask = 0, pocketMask = 0, statMask = 0, dynPairMask = 0;
                if (e < nEnt) {
                    const uint32_t ent = t.entries[e];
                    const int slot = (int)(ent >> 5), b = (int)(ent & 31u);
                    const int g = b * S + slot;
                    const float2 q0 = t.pos[g], q1 = t.vel[g], q2n = t.vn[g];
                    const int lastRail = meta_last_rail(t.meta[g]);
                    const uint32_t restMask = t.restMask[slot], dynMask = t.dynMask[slot];
                    const v2 pos = mk(q0.x, q0.y), vel = mk(q1.x, q1.y), dn = mk(q2n.x, q2n.y);
                    // dynamic partners o > b (one job covers both orders)
                    for (int o = sub; o < NB; o += TILE_K)
                        if (o > b && ((dynMask >> o) & 1u)) dynPairMask |= 1u << o;
                    if (!is_zero(vel)) {
                        for (int r = sub; r < R; r += TILE_K) {   // halfLineIntersectsLineSegment (:725-753)
                            const RailGeom rg = sRails[r];
                            const float detDV = vel.x * rg.dy - vel.y * rg.dx;
                            const float detVP = rg.dx * pos.y - rg.dy * pos.x;
                            const float tt = (detVP + sDetQV[r]) / detDV;
                            const bool useX = rg.dx != 0.0f;
                            const float num = useX ? (pos.x + tt * vel.x - rg.sx) : (pos.y + tt * vel.y - rg.sy);
                            const float u = num / (useX ? rg.dx : rg.dy);
                            if (r != lastRail && detDV != 0.0f && !(tt < 0.0f) && !(u < 0.0f || u > 1.0f)) railMask |= 1u << r;
                        }
                        for (int k = sub; k < P; k += TILE_K) {   // discriminant of lineIntersectsCircle (:584-625)
                            const PocketGeom pg = sPockets[k];
                            const v2 delta = pos - mk(pg.x, pg.y);
                            const float qa = 1;
                            const float qb = 2 * dot(dn, delta);
                            const float qc = dot(delta, delta) - pg.r * pg.r;
                            const float disc = qb * qb - 4 * qa * qc;
                            if (!(disc < 0)) pocketMask |= 1u << k;
                        }
                        const float len1 = length(vel);
                        for (int o = sub; o < NB; o += TILE_K) {  // moving vs resting (:318-324, :54-70)
                            if (!((restMask >> o) & 1u)) continue;
                            const float2 q2 = t.pos[o * S + slot];
                            const v2 toPoint = mk(q2.x, q2.y) - pos;
                            const float factor = dot(dn, toPoint) / len1;
                            const v2 dv = toPoint - (vel * factor);
                            const float dist = factor < 0 ? dot(toPoint, toPoint) : dot(dv, dv);
                            if (!(dist > dd)) statMask |= 1u << o;
                        }
                    }
                }
                push_mask(railMask, (uint32_t)ce << 5, t.qRails, &cnt.rails);
                push_mask(pocketMask, (uint32_t)ce << 5, t.qPockets, &cnt.pockets);
                push_mask(dynPairMask, (uint32_t)ce << 5, t.qDyn, &cnt.dynPairs);
                push_mask(statMask, (uint32_t)e << 5, t.qPairs, &cnt.pairs);
            }
            __syncthreads();

            // stage 1b + 2 on the chunk's queues: [rails | pockets | moving-vs-moving pre-checks], one item per thread
            {
                const int cR = cnt.rails, cP = cnt.pockets, cD = cnt.dynPairs;
                const int totalQ = cR + cP + cD;
                for (int i0 = 0; i0 < totalQ; i0 += TILE_THREADS) {
                    const int i = i0 + tid;
                    bool pass = false;
                    uint32_t pairCode = 0;
                    if (i < cR) {
                        const uint32_t c2 = t.qRails[i];
                        const uint32_t ent = t.entries[c0 + (c2 >> 5)];
                        const int r = (int)(c2 & 31u), b = (int)(ent & 31u), slot = (int)(ent >> 5);
                        const int g = b * S + slot;
                        const float2 q0 = t.pos[g], q1 = t.vel[g], q2n = t.vn[g];
                        const v2 pos = mk(q0.x, q0.y), vel = mk(q1.x, q1.y);
                        const v2 acc = ball_acc(mk(q2n.x, q2n.y), meta_rolling(t.meta[g]), pc);
                        const RailGeom rg = sRails[r];
                        // lambda of halfLineIntersectsLineSegment, recomputed exactly as stage 1 did; then :181-238
                        const float detDV = vel.x * rg.dy - vel.y * rg.dx;
                        const float detVP = rg.dx * pos.y - rg.dy * pos.x;
                        const float lambda = (detVP + sDetQV[r]) / detDV;
                        const float distance = length(lambda * vel);
                        const float velocity = length(vel);
                        const float a = -length(acc);
                        float x[2];
                        if (solve_quadratic(0.5f * a, velocity, -distance, x) != 0) offer(keyCur + slot, x[0], make_code(b, CAT_RAIL, r));
                    } else if (i < cR + cP) {
                        const uint32_t c2 = t.qPockets[i - cR];
                        const uint32_t ent = t.entries[c0 + (c2 >> 5)];
                        const int k = (int)(c2 & 31u), b = (int)(ent & 31u), slot = (int)(ent >> 5);
                        const int g = b * S + slot;
                        const float2 q0 = t.pos[g], q1 = t.vel[g], q2n = t.vn[g];
                        const v2 vn = mk(q2n.x, q2n.y);
                        const v2 acc = ball_acc(vn, meta_rolling(t.meta[g]), pc);
                        float tp;
                        if (time_to_pocket(sPockets[k], mk(q0.x, q0.y), vn, mk(q1.x, q1.y), acc, &tp))   // :383-457
                            offer(keyCur + slot, tp, make_code(b, CAT_POCKET, k));
                    } else if (i < totalQ) {
                        const uint32_t c2 = t.qDyn[i - cR - cP];
                        const int e = c0 + (int)(c2 >> 5);
                        const uint32_t ent = t.entries[e];
                        const int o = (int)(c2 & 31u), b = (int)(ent & 31u), slot = (int)(ent >> 5);
                        const int g1 = b * S + slot, g2 = o * S + slot;
                        const float2 x1 = t.pos[g1], y1 = t.vel[g1], x2 = t.pos[g2], y2 = t.vel[g2];
                        const v2 p1 = mk(x1.x, x1.y), v1 = mk(y1.x, y1.y), p2 = mk(x2.x, x2.y), w2 = mk(y2.x, y2.y);
                        // the reference tests b -> o and, on o's own turn, o -> b (:292-317); the quantised coefficients of
                        // both orders are identical, so one job covers both
                        #pragma unroll 1
                        for (int turn = 0; turn < 2 && !pass; ++turn)   // one call site, operands swapped by select
                            pass = collision_precheck(turn ? w2 : v1, turn ? p2 : p1, turn ? v1 : w2, turn ? p1 : p2, diameter);
                        pairCode = ((uint32_t)e << 5) | (uint32_t)o;
                    }
                    const unsigned m = __ballot_sync(FULL, pass);
                    if (m) {
                        int at = 0;
                        if (lane == 0) at = atomicAdd(&cnt.pairs, __popc(m));
                        at = __shfl_sync(FULL, at, 0);
                        if (pass) t.qPairs[at + __popc(m & ((1u << lane) - 1u))] = (uint16_t)pairCode;
                    }
                }
            }
            __syncthreads();
        }

        // ---- (C) the quartic: one surviving ball pair per thread ---------------------------------------------------------
        {
            const int np = cnt.pairs;
            for (int i = tid; i < np; i += TILE_THREADS) solve_pair(t, keyCur, t.qPairs[i], diameter, pc);
            if (tid == 0) {
                cQuartics += np;
                cPairs += (unsigned long long)nEnt * NB; cRails += (unsigned long long)nEnt * R; cPockets += (unsigned long long)nEnt * P;
            }
        }
        __syncthreads();
        if (tid == 0) cnt.pairs = 0;

        __syncthreads();

        // ---- (D1) every dynamic ball the event does not touch: closed-form advance (createNodes :1907-1955) --------------
        for (int i = tid; i < nEnt; i += TILE_THREADS) {
            const uint32_t ent = t.entries[i];
            const int slot = (int)(ent >> 5), b = (int)(ent & 31u);
            const unsigned long long key = keyCur[slot];
            if (key == ~0ull || t.nEvents[slot] + 1 > MAX_EVENTS) continue;   // the strike ends as it is
            const uint32_t code = (uint32_t)key;
            const int evBall = (int)(code >> 16);
            const uint32_t cat = (code >> 8) & 0xffu;
            const int evPartner = (int)(code & 0xffu);
            const bool pairEvent = cat == CAT_STATIC_PAIR || cat == CAT_DYNAMIC_PAIR;
            if (b == evBall || (pairEvent && b == evPartner)) continue;
            const float dt = from_ordered_bits((uint32_t)(key >> 32));
            const float T = t.T[slot];
            const float newT = T + dt;
            const float deltaTime = newT - T;  // type.cpp:475
            const int g = b * S + slot;
            const float2 q0 = t.pos[g], q1 = t.vel[g], q2n = t.vn[g];
            const uint32_t m = t.meta[g];
            const bool rolling = meta_rolling(m);
            v2 vn = mk(q2n.x, q2n.y), acc = ball_acc(vn, rolling, pc);
            const v2 vel0 = mk(q1.x, q1.y);
            const v2 vel = advance_velocity(acc, vel0, dt);
            const v2 pos = advance_position(acc, vel0, dt, mk(q0.x, q0.y));
            derive_ball(vel, rolling, pc, &vn, &acc);
            float nr = t.nextRolling[g];
            if (nr >= 0.0f) { const float x = nr - deltaTime; nr = x > 0.0f ? x : 0.0f; }
            if (rolling) nr = -1.0f;   // K_MOVING (type.cpp:514-518)
            t.pos[g] = make_float2(pos.x, pos.y);
            t.vel[g] = make_float2(vel.x, vel.y);
            t.vn[g] = make_float2(vn.x, vn.y);
            t.nextRolling[g] = nr;
            t.meta[g] = make_meta(K_MOVING, rolling, meta_last_rail(m), meta_type(m));
            offer(keyNext + slot, time_to_stop(acc, vel), make_code(b, CAT_REST, 0));
            if (nr >= 0.0f) offer(keyNext + slot, nr, make_code(b, CAT_ROLLING, 0));
        }
        __syncthreads();

        // ---- (D2) the winning event of every live slot (createLayer :1957-1992, System::append type.cpp:467-531) -------
        if (tid < S && t.cand[tid] >= 0) {
            const int slot = tid;
            const unsigned long long key = keyCur[slot];
            const int nEvents = t.nEvents[slot];
            if (key == ~0ull) {
                t.fin[slot] = F_NOEVENT;                       // billiard_search.cpp:1464-1467
            } else if (nEvents + 1 > MAX_EVENTS) {
                t.fin[slot] = F_OVERFLOW;                      // :1376-1379
            } else {
                const float dt = from_ordered_bits((uint32_t)(key >> 32));
                const uint32_t code = (uint32_t)key;
                const int evBall = (int)(code >> 16);
                const uint32_t cat = (code >> 8) & 0xffu;
                const int evPartner = (int)(code & 0xffu);
                const int evType = cat == CAT_REST ? E_REST : cat == CAT_ROLLING ? E_ROLLING : cat == CAT_RAIL ? E_RAIL_COLLISION
                                 : cat == CAT_POCKET ? E_POTTING : E_BALL_COLLISION;
                const float T = t.T[slot];
                const float newT = T + dt;
                const float deltaTime = newT - T;  // type.cpp:475
                uint32_t dynMask = t.dynMask[slot], restMask = t.restMask[slot];
                const int gA = evBall * S + slot;
                v2 posA, velA, vnA;
                { const float2 a = t.pos[gA], b2 = t.vel[gA], c2 = t.vn[gA]; posA = mk(a.x, a.y); velA = mk(b2.x, b2.y); vnA = mk(c2.x, c2.y); }
                const uint32_t mA = t.meta[gA];
                bool rollingA = meta_rolling(mA);
                int kindA, lastRailA = meta_last_rail(mA);
                float nrA = t.nextRolling[gA];
                const v2 accA = ball_acc(vnA, rollingA, pc);
                const v2 npos = advance_position(accA, velA, dt, posA);
                const v2 nvel = advance_velocity(accA, velA, dt);
                int cause = -1;
                posA = npos;
                if (evType == E_POTTING) {                 // :1716-1746
                    velA = mk(0.f, 0.f); rollingA = false; kindA = K_POTTING;
                } else if (evType == E_RAIL_COLLISION) {   // :1748-1786
                    const RailGeom rg = sRails[evPartner];
                    velA = rail_collision(nvel, mk(rg.nx, rg.ny), pc);
                    rollingA = true; kindA = K_RAIL; cause = evPartner;
                } else if (evType == E_REST) {             // :1788-1808
                    velA = mk(0.f, 0.f); rollingA = false; kindA = K_REST;
                } else if (evType == E_ROLLING) {          // :1810-1840
                    velA = nvel; rollingA = true; kindA = K_MOVING;
                } else {                                   // :1842-1905 createBallCollisionNodes
                    const int gB = evPartner * S + slot;
                    v2 posB, velB, vnB;
                    { const float2 a = t.pos[gB], b2 = t.vel[gB], c2 = t.vn[gB]; posB = mk(a.x, a.y); velB = mk(b2.x, b2.y); vnB = mk(c2.x, c2.y); }
                    const uint32_t mB = t.meta[gB];
                    const bool rollingB = meta_rolling(mB);
                    float nrB = t.nextRolling[gB];
                    const v2 accB = ball_acc(vnB, rollingB, pc);
                    const v2 opos = advance_position(accB, velB, dt, posB);
                    const v2 ovel = advance_velocity(accB, velB, dt);
                    v2 o1, o2;
                    elastic_collision(npos, nvel, opos, ovel, rollingA, rollingB, pc, &o1, &o2);
                    velA = o1; kindA = K_COLLISION;
                    velB = o2;
                    v2 accB2;
                    derive_ball(velB, rollingB, pc, &vnB, &accB2);
                    // System::append for the partner: K_COLLISION
                    if (nrB >= 0.0f) { const float nr = nrB - deltaTime; nrB = nr > 0.0f ? nr : 0.0f; }
                    nrB = (!rollingB && !is_zero(velB)) ? time_to_rolling(velB, pc) : -1.0f;
                    t.pos[gB] = make_float2(opos.x, opos.y);
                    t.vel[gB] = make_float2(velB.x, velB.y);
                    t.vn[gB] = make_float2(vnB.x, vnB.y);
                    t.nextRolling[gB] = nrB;
                    t.meta[gB] = make_meta(K_COLLISION, rollingB, -1, meta_type(mB));
                    offer(keyNext + slot, time_to_stop(accB2, velB), make_code(evPartner, CAT_REST, 0));
                    if (nrB >= 0.0f) offer(keyNext + slot, nrB, make_code(evPartner, CAT_ROLLING, 0));
                    dynMask |= 1u << evPartner; restMask &= ~(1u << evPartner);
                }
                v2 accA2;
                derive_ball(velA, rollingA, pc, &vnA, &accA2);
                // System::append bookkeeping (type.cpp:478-527)
                if (nrA >= 0.0f) { const float nr = nrA - deltaTime; nrA = nr > 0.0f ? nr : 0.0f; }
                if (kindA == K_COLLISION) {
                    lastRailA = -1;
                    nrA = (!rollingA && !is_zero(velA)) ? time_to_rolling(velA, pc) : -1.0f;
                } else if (kindA == K_REST || kindA == K_POTTING) {
                    nrA = -1.0f;
                } else if (kindA == K_MOVING) {
                    if (rollingA) nrA = -1.0f;
                } else if (kindA == K_RAIL) {
                    lastRailA = cause; nrA = -1.0f;
                }
                t.pos[gA] = make_float2(posA.x, posA.y);
                t.vel[gA] = make_float2(velA.x, velA.y);
                t.vn[gA] = make_float2(vnA.x, vnA.y);
                t.nextRolling[gA] = nrA;
                t.meta[gA] = make_meta(kindA, rollingA, lastRailA, meta_type(mA));
                if (kind_dynamic(kindA)) {
                    offer(keyNext + slot, time_to_stop(accA2, velA), make_code(evBall, CAT_REST, 0));
                    if (nrA >= 0.0f) offer(keyNext + slot, nrA, make_code(evBall, CAT_ROLLING, 0));
                } else {
                    dynMask &= ~(1u << evBall);
                    if (kindA == K_REST) restMask |= 1u << evBall;
                }
                t.dynMask[slot] = dynMask; t.restMask[slot] = restMask;
                t.T[slot] = newT; t.nEvents[slot] = nEvents + 1;
                if (p.out.eventLog) {
                    int a = evBall, b = evPartner;
                    if (evType == E_BALL_COLLISION && a > b) { const int x = a; a = b; b = x; }
                    if (evType == E_REST || evType == E_ROLLING) b = -1;
                    p.out.eventLog[(int64_t)t.cand[slot] * 32 + nEvents] = make_int4(evType, a, b, __float_as_int(newT));
                }
                if (dynMask == 0) t.fin[slot] = F_COMPLETED;
                cEvents += 1;   // per thread; reduced at the end
            }
        }
        __syncthreads();

        // ---- (E) finished strikes: rules, cost, results; their slots become free -----------------------------------------
        if (tid < S) {
            if (t.cand[tid] >= 0 && t.fin[tid] != 0) {
                finish_slot(p, t, tid);
                t.cand[tid] = -1; t.fin[tid] = 0;
                keyNext[tid] = ~0ull;
            }
            keyCur[tid] = ~0ull;   // becomes the next round's "next" buffer
        }
        cur ^= 1;
        __syncthreads();
    }

    if (p.out.counters) {
        if (cEvents) atomicAdd(&p.out.counters[0], cEvents);
        if (tid == 0) {
            atomicAdd(&p.out.counters[1], cPairs); atomicAdd(&p.out.counters[2], cQuartics);
            atomicAdd(&p.out.counters[3], cRails); atomicAdd(&p.out.counters[4], cPockets);
        }
    }
}

int g_tileRails = MAX_RAILS, g_tilePockets = MAX_POCKETS;

}  // namespace

void tile_upload_constants(const TableGeom& table, const RulesData& rules, cudaStream_t stream) {
    PhysConst pc = phys_const();
    cudaMemcpyToSymbolAsync(t_table, &table, sizeof(TableGeom), 0, cudaMemcpyHostToDevice, stream);
    cudaMemcpyToSymbolAsync(t_rules, &rules, sizeof(RulesData), 0, cudaMemcpyHostToDevice, stream);
    cudaMemcpyToSymbolAsync(t_phys, &pc, sizeof(PhysConst), 0, cudaMemcpyHostToDevice, stream);
    g_tileRails = table.nRails; g_tilePockets = table.nPockets;
    cudaFuncSetAttribute(tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)tile_smem_bytes(TILE_SLOTS, MAX_BALLS, MAX_RAILS, MAX_POCKETS));
}

// One launch of the batch path: `dParams` is a device copy of the launch parameters, `nextCand` a device counter.
cudaError_t launch_tile(const LaunchParams* dParams, int64_t nCand, int nBalls, int sms, unsigned long long* nextCand,
                        cudaStream_t stream) {
    if (nCand <= 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(nextCand, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    const size_t smem = tile_smem_bytes(TILE_SLOTS, nBalls, g_tileRails, g_tilePockets);
    int perSm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, tile_kernel, TILE_THREADS, smem);
    if (perSm < 1) perSm = 1;
    const int64_t need = (nCand + TILE_SLOTS - 1) / TILE_SLOTS;
    const int64_t resident = (int64_t)sms * perSm;
    tile_kernel<<<(unsigned)(need < resident ? need : resident), TILE_THREADS, smem, stream>>>(dParams, nextCand, TILE_SLOTS);
    return cudaGetLastError();
}

}  // namespace bcu
