// items_kernel of the strike pipeline (included by bcu_pipeline.cu inside its anonymous namespace).
//
// One thread per DYNAMIC BALL of the round (an entry of the stream-compacted list).  The thread loads its ball once and
// walks the R rails, P pockets and NB partner balls with the cheap first-stage tests only, collecting the survivors as
// bit masks.  Survivors of the whole CTA are then compacted into shared-memory queues and the expensive second stages
// (rail / pocket quadratics with their double-precision sqrt, the generic moving-vs-moving pre-check, the quantised
// quartic coefficients) run one queue entry per thread, i.e. on dense lanes.
//
//   rails    stage 1  halfLineIntersectsLineSegment (billiard_physics.cpp:725-753)   stage 2  the quadratic (:181-238)
//   pockets  stage 1  discriminant of lineIntersectsCircle (:584-625)                 stage 2  timeToTarget (:383-457)
//   pairs    stage 1  moving-vs-static pre-check (:318-324, pointToHalfLineSquaredDistance :54-70) with the ball's
//                     normalised velocity and speed hoisted out of the partner loop (same operations, same bits)
//            stage 1b moving-vs-moving pre-checks (:292-317), dense
//            stage 2  quantised coefficients (:326-337) -> quartic job in the global queue
#pragma once

constexpr int ITEM_BLOCK = 128;

struct ItemCounts {
    int rails, pockets, pairs, dynPairs;
};

// Appends the set bits of `mask` (as base | bit index) to a shared-memory queue; all 32 lanes must call.
// Queue entries are 16 bit: thread-in-tile << 5 | item (pairs: thread << 6 | partner is dynamic << 5 | partner); the
// thread's list entry (strike << 5 | ball) is looked up in sEnt[].
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

__global__ void __launch_bounds__(ITEM_BLOCK) items_kernel(const LaunchParams* __restrict__ pp, PipeWorkspace w, int round) {
    const LaunchParams& p = *pp;
    extern __shared__ uint16_t smemQ[];
    __shared__ ItemCounts cnt;
    __shared__ uint32_t sEnt[ITEM_BLOCK];
    __shared__ uint32_t sJobBase;
    __shared__ RailGeom sRails[MAX_RAILS];
    __shared__ float sDetQV[MAX_RAILS];
    __shared__ PocketGeom sPockets[MAX_POCKETS];
    const unsigned FULL = 0xffffffffu;
    const int tid = threadIdx.x, lane = tid & 31;
    const int R = p.table->nRails, P = p.table->nPockets, NB = p.states.nBalls;
    const uint32_t nDyn = w.counts[4 * round + 1];
    if (nDyn == 0) return;
    for (int i = tid; i < R; i += ITEM_BLOCK) {
        const RailGeom rg = p.table->rails[i];
        sRails[i] = rg;
        sDetQV[i] = rg.sx * rg.dy - rg.sy * rg.dx;  // detQV of lineIntersectsLine (:651-723): depends on the rail only
    }
    for (int i = tid; i < P; i += ITEM_BLOCK) sPockets[i] = p.table->pockets[i];
    uint16_t* qRails = smemQ;
    uint16_t* qPockets = qRails + ITEM_BLOCK * R;
    uint16_t* qPairs = qPockets + ITEM_BLOCK * P;
    uint16_t* qDyn = qPairs + ITEM_BLOCK * NB;
    const int64_t n = w.n;
    const uint32_t* dyn = w.dyn + (size_t)(round & 1) * w.dynCap;
    unsigned long long* key = w.key + (size_t)(round % 3) * w.n;
    uint32_t* jobCount = &w.counts[4 * round + 2];
    const float diameter = p.table->diameter;
    const float dd = diameter * diameter;

    for (uint32_t base = blockIdx.x * ITEM_BLOCK; base < nDyn; base += gridDim.x * ITEM_BLOCK) {
        if (tid == 0) { cnt.rails = 0; cnt.pockets = 0; cnt.pairs = 0; cnt.dynPairs = 0; }
        __syncthreads();

        // ---- stage 1: this thread's ball against every rail, pocket and resting ball ---------------------------------
        uint32_t railMask = 0, pocketMask = 0, statMask = 0, dynPairMask = 0, ent = 0;
        const uint32_t e = base + tid;
        if (e < nDyn) {
            ent = dyn[e];
            const int64_t s = ent >> 5;
            const int b = (int)(ent & 31u);
            const int64_t g = (int64_t)b * n + s;
            const float2 q0 = w.pos[g], q1 = w.vel[g], q2n = w.vn[g];
            const int lastRail = meta_last_rail(w.meta[g]);
            const uint32_t restMask = w.restMask[s], dynMask = w.dynMask[s];
            const v2 pos = mk(q0.x, q0.y), vel = mk(q1.x, q1.y), dn = mk(q2n.x, q2n.y);
            dynPairMask = dynMask & ~((2u << b) - 1u);   // dynamic partners o > b (one job covers both orders)
            if (!is_zero(vel)) {
                for (int r = 0; r < R; ++r) {
                    const RailGeom rg = sRails[r];
                    const float detDV = vel.x * rg.dy - vel.y * rg.dx;
                    const float detVP = rg.dx * pos.y - rg.dy * pos.x;
                    const float t = (detVP + sDetQV[r]) / detDV;
                    const bool useX = rg.dx != 0.0f;
                    const float num = useX ? (pos.x + t * vel.x - rg.sx) : (pos.y + t * vel.y - rg.sy);
                    const float u = num / (useX ? rg.dx : rg.dy);
                    if (r != lastRail && detDV != 0.0f && !(t < 0.0f) && !(u < 0.0f || u > 1.0f)) railMask |= 1u << r;
                }
                for (int k = 0; k < P; ++k) {
                    const PocketGeom pg = sPockets[k];
                    const v2 delta = pos - mk(pg.x, pg.y);
                    const float qa = 1;
                    const float qb = 2 * dot(dn, delta);
                    const float qc = dot(delta, delta) - pg.r * pg.r;
                    const float disc = qb * qb - 4 * qa * qc;
                    if (!(disc < 0)) pocketMask |= 1u << k;
                }
                const float len1 = length(vel);
                const float2* posOf = w.pos + s;
                for (int o = 0; o < NB; ++o) {
                    if (!((restMask >> o) & 1u)) continue;
                    const float2 q2 = posOf[(int64_t)o * n];
                    const v2 toPoint = mk(q2.x, q2.y) - pos;
                    const float factor = dot(dn, toPoint) / len1;
                    const v2 dv = toPoint - (vel * factor);
                    const float dist = factor < 0 ? dot(toPoint, toPoint) : dot(dv, dv);
                    if (!(dist > dd)) statMask |= 1u << o;
                }
            }
        }
        sEnt[tid] = ent;
        push_mask(railMask, (uint32_t)tid << 5, qRails, &cnt.rails);
        push_mask(pocketMask, (uint32_t)tid << 5, qPockets, &cnt.pockets);
        push_mask(statMask, (uint32_t)tid << 6, qPairs, &cnt.pairs);
        push_mask(dynPairMask, (uint32_t)tid << 5, qDyn, &cnt.dynPairs);
        __syncthreads();

        // ---- stage 1b: moving-vs-moving pre-checks, one queued pair per thread ----------------------------------------
        {
            const int nd = cnt.dynPairs;
            for (int i0 = 0; i0 < nd; i0 += ITEM_BLOCK) {
                const int i = i0 + tid;
                bool pass = false;
                uint32_t code = 0;
                if (i < nd) {
                    const uint32_t c2 = qDyn[i];
                    const uint32_t en = sEnt[c2 >> 5];
                    const int o = (int)(c2 & 31u), b = (int)(en & 31u);
                    const int64_t s = en >> 5, g1 = (int64_t)b * n + s, g2 = (int64_t)o * n + s;
                    const float2 x1 = w.pos[g1], y1 = w.vel[g1], n1 = w.vn[g1], x2 = w.pos[g2], y2 = w.vel[g2], n2 = w.vn[g2];
                    const v2 p1 = mk(x1.x, x1.y), v1 = mk(y1.x, y1.y), p2 = mk(x2.x, x2.y), w2 = mk(y2.x, y2.y);
                    const v2 v1n = mk(n1.x, n1.y), w2n = mk(n2.x, n2.y);
                    // the reference tests b -> o and, on o's own turn, o -> b; the quantised coefficients of both
                    // orders are identical, so one job covers both
                    #pragma unroll 1
                    for (int turn = 0; turn < 2 && !pass; ++turn) {   // one call site each, operands swapped by select
                        const v2 va = turn ? w2 : v1, pa = turn ? p2 : p1, van = turn ? w2n : v1n;
                        const v2 vb = turn ? v1 : w2, pb = turn ? p1 : p2, vbn = turn ? v1n : w2n;
                        if (is_zero(va)) continue;                                  // :287
                        if (dot(vb, vb) > 0) pass = collision_precheck_moving(va, pa, van, vb, pb, vbn, diameter);
                        else pass = !(point_to_half_line_sq(pa, va, pb) > diameter * diameter);   // :318-324
                    }
                    code = ((c2 >> 5) << 6) | 32u | (uint32_t)o;
                }
                const unsigned m = __ballot_sync(FULL, pass);
                if (m) {
                    int at = 0;
                    if (lane == 0) at = atomicAdd(&cnt.pairs, __popc(m));
                    at = __shfl_sync(FULL, at, 0);
                    if (pass) qPairs[at + __popc(m & ((1u << lane) - 1u))] = (uint16_t)code;
                }
            }
        }
        __syncthreads();

        // ---- stage 2: one survivor per thread over the concatenated queues [rails | pockets | pairs] -------------------
        {
            const int cR = cnt.rails, cP = cnt.pockets, cQ = cnt.pairs;
            const int totalQ = cR + cP + cQ;
            if (tid == 0 && cQ > 0) sJobBase = atomicAdd(jobCount, (uint32_t)cQ);   // one global atomic per tile
            __syncthreads();
            for (int i0 = 0; i0 < totalQ; i0 += ITEM_BLOCK) {
                const int i = i0 + tid;
                const bool isRail = i < cR, isPocket = !isRail && i < cR + cP, isPair = !isRail && !isPocket && i < totalQ;
                if (isRail) {
                    const uint32_t c2 = qRails[i];
                    const uint32_t en = sEnt[c2 >> 5];
                    const int r = (int)(c2 & 31u), b = (int)(en & 31u);
                    const int64_t s = en >> 5, g = (int64_t)b * n + s;
                    const float2 q0 = w.pos[g], q1 = w.vel[g], q3 = w.acc[g];
                    const v2 pos = mk(q0.x, q0.y), vel = mk(q1.x, q1.y);
                    const RailGeom rg = sRails[r];
                    // lambda of halfLineIntersectsLineSegment, recomputed exactly as stage 1 did
                    const float detDV = vel.x * rg.dy - vel.y * rg.dx;
                    const float detVP = rg.dx * pos.y - rg.dy * pos.x;
                    const float lambda = (detVP + sDetQV[r]) / detDV;
                    const float distance = length(lambda * vel);
                    const float velocity = length(vel);
                    const float a = -length(mk(q3.x, q3.y));
                    float x[2];
                    if (solve_quadratic(0.5f * a, velocity, -distance, x) != 0) offer(key + s, x[0], make_code(b, CAT_RAIL, r));
                } else if (isPocket) {
                    const uint32_t c2 = qPockets[i - cR];
                    const uint32_t en = sEnt[c2 >> 5];
                    const int k = (int)(c2 & 31u), b = (int)(en & 31u);
                    const int64_t s = en >> 5, g = (int64_t)b * n + s;
                    const float2 q0 = w.pos[g], q1 = w.vel[g], q2n = w.vn[g], q3 = w.acc[g];
                    float t;
                    if (time_to_pocket(sPockets[k], mk(q0.x, q0.y), mk(q2n.x, q2n.y), mk(q1.x, q1.y), mk(q3.x, q3.y), &t))
                        offer(key + s, t, make_code(b, CAT_POCKET, k));
                }
                else if (isPair) {   // quantised coefficients -> global job queue
                    const int iq = i - cR - cP;
                    const uint32_t c2 = qPairs[iq];
                    const uint32_t en = sEnt[c2 >> 6];
                    const int o = (int)(c2 & 31u), b = (int)(en & 31u);
                    const int64_t s = en >> 5, g1 = (int64_t)b * n + s, g2 = (int64_t)o * n + s;
                    const float2 x1 = w.pos[g1], y1 = w.vel[g1], z1 = w.acc[g1], x2 = w.pos[g2], y2 = w.vel[g2], z2 = w.acc[g2];
                    float co[5];
                    collision_coefficients(mk(z1.x, z1.y), mk(y1.x, y1.y), mk(x1.x, x1.y), mk(z2.x, z2.y), mk(y2.x, y2.y),
                                           mk(x2.x, x2.y), diameter, co);
                    const uint64_t j = (uint64_t)sJobBase + iq;
                    if (j < w.jobCap) {
                        #pragma unroll
                        for (int q = 0; q < 5; ++q) w.jobCo[(size_t)q * w.jobCap + j] = co[q];
                        w.jobCode[j] = (en << 6) | (c2 & 63u);
                    } else {
                        w.counts[4 * ROUNDS + 3] = 1u;  // queue overflow: the host grows the queue and re-runs the launch
                    }
                }
            }
        }
        __syncthreads();
    }
}
