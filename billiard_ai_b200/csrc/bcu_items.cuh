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

#ifndef ITEM_BLOCK_V
#define ITEM_BLOCK_V 128
#endif
constexpr int ITEM_BLOCK = ITEM_BLOCK_V;
#ifndef ITEM_MIN_BLOCKS
#define ITEM_MIN_BLOCKS 6
#endif
#ifndef ITEM_DENSE_MIN_BLOCKS
#define ITEM_DENSE_MIN_BLOCKS 8   // items_kernel_dense: the instantiation for large parts (see launch_pipeline)
#endif

struct ItemCounts {
    int pairs, dynPairs;
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

// ---- the reach bound --------------------------------------------------------------------------------------------------
// A ball-pair quartic is only worth solving if its result could still beat the strike's best event time so far.  The
// quartic of physics::timeToCollision (:326-337) is  |dP + dV t + dA t^2 / 2|^2 - d^2  with coefficients quantised to
// 1e-6; the reference accepts a computed root x + iy as the collision time when |y| < 1e-3 (realRoot :550-560) and takes
// the smallest such x, or none if that one is negative.  For any such root of the polynomial with 0 <= x <= T:
//
//     |dP| <= d + disp1(T) + disp2(T) + 1e-3 (|v1| + |v2| + (|a1| + |a2|) T) + q(T)
//
// where disp(T) = max(|v| T, |a| T^2 / 2) bounds |v t + a t^2 / 2| on [0, T] for an acceleration anti-parallel to the
// velocity (the polynomial keeps decelerating past the stop, so the ball "comes back"), the third term is the change of
// the polynomial over the 1e-3 imaginary strip, and q(T) = 0.05 mm (1 + T + T^2 + T^3 + T^4) covers the quantisation and
// binary32 rounding of the coefficients.  A pair whose distance exceeds this REACH for T = (best time so far) cannot
// produce the winning event -- PROVIDED the closed-form solver returns the polynomial's roots.  It does not always: when
// the relative motion runs along a line (every moving-vs-resting pair: dA || dV) the depressed quartic is biquadratic up to
// quantisation noise, Ferrari's S is then ~0 unless all four roots are real, and q/S is rounding noise that the reference
// turns into an "event" about once in 5e5 such pairs.  Those results are reproduced bit for bit by solving the quartic;
// the bound is only applied where the solver resolves the roots:
//   * moving vs resting: only if the ball's polynomial passes the partner completely before its turning point
//     (|dP| + d + 5 mm <= 0.99 v^2 / 2|a|: four real roots, S = half the sum of the two larger ones); every other resting
//     partner is solved (a partner beyond the turning point has four complex roots and "should" give no event, but that is
//     exactly where the closed form is noise: DESIGN.md section 4);
//   * moving vs moving: only if dA and dV are more than 0.57 degrees apart (sin^2 >= 1e-4: q is then 4 orders of magnitude
//     above the quantisation noise that makes the collinear case fragile) and the leading coefficient is >= 1e-5 (a = 0
//     exactly makes the reference return "none" before solving; in between the quartic is nearly cubic).  Audits with
//     both guards removed found no failure in 5.4e8 marked jobs, but a later one (5.2e7 strikes, 16 balls) did find ONE:
//     two rolling balls 1.08 m apart whose quantised polynomial has the roots 2.05 +- 124.7i, 2.05 +- 1.80i -- all four with
//     the same real part, i.e. the depressed quartic's q = (b^3 - 4abc + 8a^2 d) / 8a^3 cancels to 1e-3 of its terms by
//     accident (no collinearity: sin^2 = 0.999), S ~ 0 again, and the reference reports a "collision" at 0.108 s.  So the
//     condition that matters is the one on q itself: the bound is used only if |b^3 - 4abc + 8a^2 d| >= 1e-2 x the sum of the
//     magnitudes of its three terms, evaluated on the quantised binary32 coefficients the solver will see (this contains
//     the collinear case; the two older guards stay as margin).  Measured on that polynomial with d swept through the
//     cancellation (the reference's own solver against numpy's roots): ratio 4e-6 (the failure) garbage, 5e-5 -> roots off
//     by 3e-3 s, 5e-4 -> 1e-4 s, 5e-3 -> 4e-7 s, 5e-2 -> 6e-9 s; the bound's own slack is 2e-3 s.
// On top of the bound 2 ms and 2 mm of slack are left for the rounding of the solver in these regimes.  The strike's key
// only decreases during a round, so any value read from it is a valid T.  mode 2 (audit) posts everything and marks the
// jobs the bound would have skipped; quartic_kernel then counts the marked jobs whose result would have mattered
// (tools/prune_audit.py, tests/test_gpu_parity.py).
struct Reach {
    float t;       // T + slack, seconds (NaN: no bound)
    float fixed;   // d + 2 mm + q(t)
};
__device__ __forceinline__ Reach reach_of(float T, float diameter) {
    Reach r;
    r.t = T + 2e-3f;
    const float poly = (((r.t + 1.0f) * r.t + 1.0f) * r.t + 1.0f) * r.t + 1.0f;
    r.fixed = diameter + 2.0f + 0.05f * poly;
    return r;
}
// reach contribution of one ball with speed v and deceleration a (both >= 0)
__device__ __forceinline__ float reach_ball(const Reach& r, float v, float a) {
    const float lin = v * r.t, quad = 0.5f * a * r.t * r.t;
    return (lin > quad ? lin : quad) + 1e-3f * (v + a * r.t);
}
#ifndef PRUNE_MIN_LEADING_V
#define PRUNE_MIN_LEADING_V 1e-5f
#endif
#ifndef PRUNE_MIN_SIN_SQ_V
#define PRUNE_MIN_SIN_SQ_V 1e-4f
#endif
#ifndef PRUNE_MIN_QREL_V
#define PRUNE_MIN_QREL_V 1e-2f
#endif
constexpr float PRUNE_MIN_QREL = PRUNE_MIN_QREL_V;   // |depressed q's numerator| / (sum of |its terms|), see above
constexpr float PRUNE_MIN_LEADING = PRUNE_MIN_LEADING_V;   // m^2/s^4; moving-vs-resting pairs have 4.87e-3 (rolling) or 0.70 (sliding)
constexpr float PRUNE_MIN_SIN_SQ = PRUNE_MIN_SIN_SQ_V;
constexpr uint32_t JOB_MARK = 0x80000000u;
#ifndef PRUNE_REST_PASS_ONLY_V
#define PRUNE_REST_PASS_ONLY_V 1   // 0: apply the bound to every resting partner (a weakened build the regression vectors must catch)
#endif
#ifndef PARTNER_BATCH_V
#define PARTNER_BATCH_V 11
#endif
constexpr int PARTNER_BATCH = PARTNER_BATCH_V;   // resting partners loaded per memory round trip (22 balls: two trips)   // audit: the bound would have skipped this job

#ifndef JOB_SPLIT
#define JOB_SPLIT 1
#endif
// Which arm of the closed-form solver a job will take, read off its coefficients: std::pow(complex, double) on the base
// (deltaOne + sqrt(deltaOne^2 - 4 deltaZero^3)) / 2 (solve_quartic_roots) is a real pow() when that base is a positive real --
// R = deltaOne^2 - 4 deltaZero^3 >= 0 and deltaOne + sqrt(R) > 0, i.e. deltaOne > 0 or deltaZero < 0 -- and clog + exp + sincos
// otherwise, and the complex square roots and divisions downstream follow the same split (real against complex values).
// A scheduling hint only: evaluated in plain double, and a wrong guess next to a boundary costs divergence, nothing else.
__device__ __forceinline__ bool job_class_real(const float co[5]) {
    const double a = co[0], b = co[1], c = co[2], d = co[3], e = co[4];
    const double dz = (c * c) - (3.0 * b * d) + (12.0 * a * e);
    const double d1 = (2.0 * c * c * c) - (9.0 * b * c * d) + (27.0 * b * b * e) + (27.0 * a * d * d) - (72.0 * a * c * e);
    const double R = d1 * d1 - 4.0 * dz * dz * dz;
    return R >= 0.0 && (d1 > 0.0 || dz < 0.0);
}

// shared memory of the items stage (one instance per CTA)
struct ItemShared {
    ItemCounts cnt;
    uint32_t jobBase;
    uint32_t splitBase[2][2];                    // JOB_SPLIT: [trip parity][class] first slot of this trip's jobs of a class
    uint32_t splitCnt[2][ITEM_BLOCK / 32][2];    // JOB_SPLIT: [trip parity][warp][class] jobs of a class in a warp
    uint32_t ent[ITEM_BLOCK];
    float T[ITEM_BLOCK];
    RailGeom rails[MAX_RAILS];
    float detQV[MAX_RAILS];
    float railLenSq[MAX_RAILS];   // |end - start|^2, or -1 when the component u is divided by is too short to filter on
    PocketGeom pockets[MAX_POCKETS];
};

// table geometry -> shared memory; depends on nothing a kernel of the launch writes
__device__ __forceinline__ void items_prologue(ItemShared& sh, const PipeWorkspace& w) {
    const int tid = threadIdx.x;
    const int R = w.nRails, P = w.nPockets;
    for (int i = tid; i < R; i += ITEM_BLOCK) {
        const RailGeom rg = w.table->rails[i];
        sh.rails[i] = rg;
        sh.detQV[i] = rg.sx * rg.dy - rg.sy * rg.dx;  // detQV of lineIntersectsLine (:651-723): depends on the rail only
        const float used = fabsf(rg.dx != 0.0f ? rg.dx : rg.dy);
        sh.railLenSq[i] = used >= 5.0f ? rg.dx * rg.dx + rg.dy * rg.dy : -1.0f;
    }
    for (int i = tid; i < P; i += ITEM_BLOCK) sh.pockets[i] = w.table->pockets[i];
}

// `bid` of `nb` CTAs work on this round (items_kernel passes its grid)
__device__ __forceinline__ void items_body(ItemShared& sh, uint16_t* smemQ, const LaunchParams& p, const PipeWorkspace& w, int round,
                                           unsigned bid, unsigned nb) {
    ItemCounts& cnt = sh.cnt;
    uint32_t* sEnt = sh.ent;
    float* sT = sh.T;
    uint32_t& sJobBase = sh.jobBase;
    const RailGeom* sRails = sh.rails;
    const float* sDetQV = sh.detQV;
    const float* sRailLenSq = sh.railLenSq;
    const PocketGeom* sPockets = sh.pockets;
    const unsigned FULL = 0xffffffffu;
    const int tid = threadIdx.x, lane = tid & 31;
    const int R = w.nRails, P = w.nPockets, NB = w.nBalls;
    const uint32_t nDyn = w.counts[4 * round + 1];
    if (nDyn == 0) return;
    uint16_t* qPairs = smemQ;
    uint16_t* qDyn = qPairs + ITEM_BLOCK * NB;
    const int64_t n = w.n;
    const uint32_t* dyn = w.dyn + (size_t)(round & 1) * w.dynCap;
    unsigned long long* key = w.key + (size_t)(round % 3) * w.n;
    uint32_t* jobCount = &w.counts[4 * round + 2];
    const float diameter = w.diameter;
    const float dd = diameter * diameter;
    const int prune = w.pruneMode;
    unsigned long long skipped = 0;   // counters / audit only

    for (uint32_t base = bid * ITEM_BLOCK; base < nDyn; base += nb * ITEM_BLOCK) {
        if (tid == 0) { cnt.pairs = 0; cnt.dynPairs = 0; }
        __syncthreads();

        // ---- stage 1: this thread's ball against every rail, pocket and resting ball ---------------------------------
        uint32_t statMask = 0, statMarked = 0, dynPairMask = 0, ent = 0;
        float tBest = __int_as_float(0x7fc00000);
        const uint32_t e = base + tid;
        if (e < nDyn) {
            ent = dyn[e];
            const int64_t s = ent >> 5;
            const int b = (int)(ent & 31u);
            const int64_t g = (int64_t)b * n + s;
            const float2 q0 = w.pos[g], q1 = w.vel[g], q2n = w.vn[g], q3 = w.acc[g];
            const uint32_t metaWord = w.meta[g];
            const uint32_t lineWord = w.lineCand[g];   // loaded with the state (used if the ball kept its line): no extra round trip
            const int lastRail = meta_last_rail(metaWord);
            const uint32_t restMask = w.restMask[s], dynMask = w.dynMask[s];
            unsigned long long keyNow = __ldcg(key + s);   // rest / rolling offers of the strike (and whatever else is in)
            const v2 pos = mk(q0.x, q0.y), vel = mk(q1.x, q1.y), dn = mk(q2n.x, q2n.y), acc = mk(q3.x, q3.y);
            dynPairMask = dynMask & ~((2u << b) - 1u);   // dynamic partners o > b (one job covers both orders)
            if (!is_zero(vel)) {
                uint32_t railMask = 0, pocketMask = 0;
                // The rail and pocket candidates are functions of the ball's LINE of motion (plus "already behind", which only
                // grows): a ball that was merely advanced since the last round (kind MOVING: same direction, further along)
                // reuses the masks stored when its direction was set; the exact stages below sort out what it has passed since.
                const bool reuse = prune && meta_kind(metaWord) == K_MOVING && R <= 24;
                if (reuse) {
                    const uint32_t c2 = lineWord;
                    railMask = c2 & 0xffffffu;
                    pocketMask = c2 >> 24;
                } else {
                    // Rails, pass A (no division): which rails can the half-line cross at all?  The exact test below divides twice
                    // per rail; here a rail is dropped only when the exact test is certain to fail: the intersection parameter
                    // t = N / detDV is negative by more than any rounding (N detDV < 0, |t| > 1e-10), or -- for lines at more than
                    // ~1.1 degrees (sin^2 >= 4e-4), where both this estimate and the reference's u are accurate to < 1e-2 -- the
                    // crossing lies more than 5 % of the segment outside it (prune mode 0: the first test only).  Everything else is a candidate.
                    const float speedSq = dot(vel, vel);
                    for (int r = 0; r < R; ++r) {
                        const RailGeom rg = sRails[r];
                        const float detDV = vel.x * rg.dy - vel.y * rg.dx;
                        const float nT = (rg.dx * pos.y - rg.dy * pos.x) + sDetQV[r];
                        // "behind" is exact, not an estimate -- nT and detDV are the very operands of the exact stage's division
                        // lambda = nT / detDV, and a quotient of opposite signs and magnitude > 1e-10 is a negative float whatever
                        // the rounding, which that stage rejects (lambda < 0) -- so it also holds with the filters off
                        const bool behind = nT * detDV < 0.0f && fabsf(nT) > 1e-10f * fabsf(detDV);
                        bool cand = r != lastRail && detDV != 0.0f && !behind;
                        if (prune) {
                            const float lenSq = sRailLenSq[r];
                            const float nU = (pos.x - rg.sx) * vel.y - (pos.y - rg.sy) * vel.x;   // u = nU / -detDV
                            const float den = -detDV, sgnU = den < 0.0f ? -nU : nU, aden = fabsf(den);
                            const bool wellCond = lenSq > 0.0f && detDV * detDV >= 4e-4f * (speedSq * lenSq);
                            const bool outside = wellCond && (sgnU < -0.05f * aden || sgnU > 1.05f * aden);
                            cand = cand && !outside;
                        }
                        if (cand) railMask |= 1u << r;
                    }
                    // Pockets, first stage: discriminant of lineIntersectsCircle (:584-625).  A pocket the ball is outside of and
                    // moving away from (qb > 0, qc > 0) has two negative intersection parameters, so timeToTarget (:383-457) returns
                    // none for it whatever the rounding (q < 0, hence x0 = q / a < 0 and x1 = c / q <= -0): dropped here.
                    for (int k = 0; k < P; ++k) {
                        const PocketGeom pg = sPockets[k];
                        const v2 delta = pos - mk(pg.x, pg.y);
                        const float qa = 1;
                        const float qb = 2 * dot(dn, delta);
                        const float qc = dot(delta, delta) - pg.r * pg.r;
                        const float disc = qb * qb - 4 * qa * qc;
                        // (with the filters on, a pocket is kept down to 0.4 mm outside the tangent: the mask outlives this position)
                        const float slack = prune ? 1e-5f * (qb * qb + 4.0f * (dot(delta, delta) + pg.r * pg.r)) + 1.0f : 0.0f;
                        if (!(disc < -slack) && !(prune && qb > 0.0f && qc > 0.0f)) pocketMask |= 1u << k;
                    }
                    if (prune && R <= 24) w.lineCand[g] = railMask | (pocketMask << 24);
                }
                const float len1 = length(vel);
                const float accLen = length(acc);
                // second stages of this ball's rails and pockets (a handful per ball): the quadratics of
                // timeToCollisionWithRail (:181-238) and timeToTarget (:383-457); the earliest goes to the strike's key
                unsigned long long best = ~0ull;
                while (railMask) {
                    const int r = __ffs(railMask) - 1;
                    railMask &= railMask - 1;
                    const RailGeom rg = sRails[r];
                    // halfLineIntersectsLineSegment (:725-753) via lineIntersectsLine (:651-723), exactly
                    const float detDV = vel.x * rg.dy - vel.y * rg.dx;
                    const float detVP = rg.dx * pos.y - rg.dy * pos.x;
                    const float lambda = (detVP + sDetQV[r]) / detDV;
                    const bool useX = rg.dx != 0.0f;
                    const float num = useX ? (pos.x + lambda * vel.x - rg.sx) : (pos.y + lambda * vel.y - rg.sy);
                    const float u = num / (useX ? rg.dx : rg.dy);
                    if (lambda < 0.0f || u < 0.0f || u > 1.0f) continue;
                    const float distance = length(lambda * vel);
                    float x[2];
                    if (solve_quadratic(0.5f * -accLen, len1, -distance, x) != 0 && x[0] == x[0]) {
                        const unsigned long long k2 = ((unsigned long long)ordered_bits(x[0]) << 32) | make_code(b, CAT_RAIL, r);
                        best = k2 < best ? k2 : best;
                    }
                }
                while (pocketMask) {
                    const int k = __ffs(pocketMask) - 1;
                    pocketMask &= pocketMask - 1;
                    if (reuse) {   // passed since the mask was made? (outside and moving away: none, see the first stage)
                        const PocketGeom pg = sPockets[k];
                        const v2 delta = pos - mk(pg.x, pg.y);
                        if (2 * dot(dn, delta) > 0.0f && dot(delta, delta) - pg.r * pg.r > 0.0f) continue;
                    }
                    float t;
                    if (time_to_pocket(sPockets[k], pos, dn, vel, acc, &t) && t == t) {
                        const unsigned long long k2 = ((unsigned long long)ordered_bits(t) << 32) | make_code(b, CAT_POCKET, k);
                        best = k2 < best ? k2 : best;
                    }
                }
                if (best != ~0ull) atomicMin(key + s, best);
                keyNow = best < keyNow ? best : keyNow;
                tBest = from_ordered_bits((uint32_t)(keyNow >> 32));   // NaN when the strike has no offer at all
                const Reach rc = reach_of(tBest, diameter);
                const float reach = rc.fixed + reach_ball(rc, len1, accLen);
                const float reachSq = reach * reach;
                // resting partners this far at most are passed completely before the turning point (see above)
                const float passMax = 0.99f * (len1 * len1) / (2.0f * accLen) - diameter - 5.0f;
                const float passMaxSq = passMax > 0.0f ? passMax * passMax : -1.0f;
                // Resting partners, pass A (no division): a partner farther than d + 0.5 mm from the LINE of motion fails the
                // reference's corridor test (:318-324) whichever arm it takes (the cross product with the stored unit velocity
                // is the perpendicular distance to ~1e-3 mm); only the others get the exact test.  Prune mode 0: all of them.
                const float2* posOf = w.pos + s;
                uint32_t candMask = 0;
                const float corridor = diameter + 0.5f;
                for (int o0 = 0; o0 < NB; o0 += PARTNER_BATCH) {
                    float2 qq[PARTNER_BATCH];
                    const float2* src = posOf + (int64_t)o0 * n;
                    #pragma unroll
                    for (int k = 0; k < PARTNER_BATCH; ++k) {   // independent loads first: one memory round trip per batch
                        qq[k] = (o0 + k < NB && ((restMask >> (o0 + k)) & 1u)) ? *src : make_float2(0.f, 0.f);
                        src += n;
                    }
                    #pragma unroll
                    for (int k = 0; k < PARTNER_BATCH; ++k) {
                        const float cr = dn.x * (qq[k].y - pos.y) - dn.y * (qq[k].x - pos.x);
                        if (!prune || !(fabsf(cr) > corridor)) candMask |= 1u << (o0 + k);
                    }
                }
                candMask &= restMask;
                while (candMask) {
                    const int o = __ffs(candMask) - 1;
                    candMask &= candMask - 1;
                    const float2 q2 = posOf[(int64_t)o * n];
                    const v2 toPoint = mk(q2.x, q2.y) - pos;
                    const float factor = dot(dn, toPoint) / len1;
                    const v2 dv = toPoint - (vel * factor);
                    const float distSq = dot(toPoint, toPoint);
                    const float dist = factor < 0 ? distSq : dot(dv, dv);
                    if (!(dist > dd)) {
                        bool skip = false;
                        if (prune) {
#if PRUNE_REST_PASS_ONLY_V
                            if (distSq <= passMaxSq) skip = distSq > reachSq;   // passed completely before the turning point
#else
                            skip = distSq > reachSq;   // guard variant for tests/golden/prune_regressions.json: NOT sound
#endif
                        }
                        if (skip) {
                            ++skipped;
                            if (prune == 2) { statMask |= 1u << o; statMarked |= 1u << o; }
                        } else {
                            statMask |= 1u << o;
                        }
                    }
                }
            }
        }
        sEnt[tid] = ent;
        sT[tid] = tBest;
        push_mask(statMask & ~statMarked, (uint32_t)tid << 6, qPairs, &cnt.pairs);
        if (prune == 2) push_mask(statMarked, ((uint32_t)tid << 6) | 0x8000u, qPairs, &cnt.pairs);
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
                    const float2 z1 = w.acc[g1], z2 = w.acc[g2];
                    const v2 p1 = mk(x1.x, x1.y), v1 = mk(y1.x, y1.y), p2 = mk(x2.x, x2.y), w2 = mk(y2.x, y2.y);
                    const v2 v1n = mk(n1.x, n1.y), w2n = mk(n2.x, n2.y);
                    // reach bound first (it needs no division): a pair that cannot meet before the strike's best time is
                    // not even pre-checked.  Audit mode pre-checks everything and marks what the bound would have dropped.
                    bool outOfReach = false;
                    if (prune) {
                        const Reach rc = reach_of(sT[c2 >> 5], diameter);
                        const v2 a1 = mk(z1.x, z1.y), a2 = mk(z2.x, z2.y);
                        const float reach = rc.fixed + reach_ball(rc, length(v1), length(a1)) + reach_ball(rc, length(w2), length(a2));
                        const v2 dA = (a1 - a2) * (1.0f / 1000.0f), dV = (v1 - w2) * (1.0f / 1000.0f), dP = p2 - p1;
                        const float aa = dot(dA, dA), vv = dot(dV, dV), cr = dA.x * dV.y - dA.y * dV.x;
                        outOfReach = 0.25f * aa >= PRUNE_MIN_LEADING && vv > 0.0f && cr * cr >= PRUNE_MIN_SIN_SQ * (aa * vv) &&
                                     !is_zero(v1) && !is_zero(w2) && dot(dP, dP) > reach * reach;
                        if (outOfReach) {   // is the solver's S resolved?  (the depressed quartic must not be nearly biquadratic)
                            float co[5];
                            collision_coefficients(a1, v1, p1, a2, w2, p2, diameter, co);
                            const float t1 = co[1] * co[1] * co[1], t2 = 4.0f * co[0] * co[1] * co[2], t3 = 8.0f * co[0] * co[0] * co[3];
                            outOfReach = fabsf((t1 - t2) + t3) >= PRUNE_MIN_QREL * (fabsf(t1) + fabsf(t2) + fabsf(t3));
                        }
                    }
                    // the reference tests b -> o and, on o's own turn, o -> b; the quantised coefficients of both
                    // orders are identical, so one job covers both
                    if (!outOfReach || prune == 2) {
                        #pragma unroll 1
                        for (int turn = 0; turn < 2 && !pass; ++turn) {   // one call site each, operands swapped by select
                            const v2 va = turn ? w2 : v1, pa = turn ? p2 : p1, van = turn ? w2n : v1n;
                            const v2 vb = turn ? v1 : w2, pb = turn ? p1 : p2, vbn = turn ? v1n : w2n;
                            if (is_zero(va)) continue;                                  // :287
                            if (dot(vb, vb) > 0) pass = collision_precheck_moving(va, pa, van, vb, pb, vbn, diameter);
                            else pass = !(point_to_half_line_sq(pa, va, pb) > diameter * diameter);   // :318-324
                        }
                    }
                    code = ((c2 >> 5) << 6) | 32u | (uint32_t)o;
                    if (pass && outOfReach) {   // audit only
                        ++skipped;
                        code |= 0x8000u;
                    } else if (outOfReach && prune != 2) {
                        ++skipped;              // mode 1: pairs dropped before the pre-check (the reference solves only those
                    }                           // that pass it: bench.py takes the reference's count from an audit launch)
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

        // ---- stage 2: quantised coefficients of the surviving pairs, one per thread -> global job queue -------------------
#if JOB_SPLIT
        // The queue is filled from both ends, by the arm of the solver a job will take (job_class_real): "real" jobs from slot
        // 0 upwards (counts[4 round + 2]), "complex" jobs from the last slot downwards (counts[4 round + 3]), so that the warps
        // of quartic_kernel -- 32 consecutive slots each -- hold jobs of ONE class (all but one warp of the grid).
        {
            const int cQ = cnt.pairs;
            int trip = 0;
            for (int iq0 = 0; iq0 < cQ; iq0 += ITEM_BLOCK, ++trip) {   // block-uniform
                const int iq = iq0 + tid, par = trip & 1;
                const bool have = iq < cQ;
                float co[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
                uint32_t jc = 0;
                bool real = false;
                if (have) {
                    const uint32_t c2 = qPairs[iq];
                    const uint32_t en = sEnt[(c2 >> 6) & (ITEM_BLOCK - 1)];
                    const int o = (int)(c2 & 31u), b = (int)(en & 31u);
                    const int64_t s = en >> 5, g1 = (int64_t)b * n + s, g2 = (int64_t)o * n + s;
                    const bool partnerDynamic = (c2 & 32u) != 0;   // a resting partner is position only (rest_ball(), bcu_pipeline.cu)
                    const float2 x1 = w.pos[g1], y1 = w.vel[g1], z1 = w.acc[g1], x2 = w.pos[g2];
                    float2 y2 = make_float2(0.f, 0.f), z2 = make_float2(0.f, 0.f);
                    if (partnerDynamic) { y2 = w.vel[g2]; z2 = w.acc[g2]; }
                    const v2 a2 = partnerDynamic ? mk(z2.x, z2.y) : rest_acc(p_phys);
                    collision_coefficients(mk(z1.x, z1.y), mk(y1.x, y1.y), mk(x1.x, x1.y), a2, mk(y2.x, y2.y),
                                           mk(x2.x, x2.y), diameter, co);
                    jc = (en << 6) | (c2 & 63u) | ((c2 & 0x8000u) ? JOB_MARK : 0u);
                    real = job_class_real(co);
                }
                const unsigned mR = __ballot_sync(FULL, have && real), mC = __ballot_sync(FULL, have && !real);
                if (lane == 0) { sh.splitCnt[par][tid >> 5][0] = __popc(mR); sh.splitCnt[par][tid >> 5][1] = __popc(mC); }
                __syncthreads();
                if (tid == 0) {
                    uint32_t tR = 0, tC = 0;
                    #pragma unroll
                    for (int v = 0; v < ITEM_BLOCK / 32; ++v) { tR += sh.splitCnt[par][v][0]; tC += sh.splitCnt[par][v][1]; }
                    if (tR) sh.splitBase[par][0] = atomicAdd(jobCount, tR);
                    if (tC) sh.splitBase[par][1] = atomicAdd(jobCount + 1, tC);
                }
                __syncthreads();
                if (have) {
                    const int k = real ? 0 : 1;
                    uint32_t at = sh.splitBase[par][k] + __popc((real ? mR : mC) & ((1u << lane) - 1u));
                    #pragma unroll
                    for (int v = 0; v < ITEM_BLOCK / 32; ++v) at += v < (tid >> 5) ? sh.splitCnt[par][v][k] : 0u;
                    if (at < w.jobCap) {
                        const uint64_t j = real ? (uint64_t)at : (uint64_t)w.jobCap - 1u - at;
                        #pragma unroll
                        for (int q = 0; q < 5; ++q) w.jobCo[(size_t)q * w.jobCap + j] = co[q];
                        w.jobCode[j] = jc;
                    } else {
                        w.counts[4 * ROUNDS + 3] = 1u;  // queue overflow: the host grows the queue and re-runs the launch
                    }
                }
            }
        }
#else
        {
            const int cQ = cnt.pairs;
            if (tid == 0 && cQ > 0) sJobBase = atomicAdd(jobCount, (uint32_t)cQ);   // one global atomic per tile
            __syncthreads();
            for (int iq = tid; iq < cQ; iq += ITEM_BLOCK) {
                const uint32_t c2 = qPairs[iq];
                const uint32_t en = sEnt[(c2 >> 6) & (ITEM_BLOCK - 1)];
                const int o = (int)(c2 & 31u), b = (int)(en & 31u);
                const int64_t s = en >> 5, g1 = (int64_t)b * n + s, g2 = (int64_t)o * n + s;
                const bool partnerDynamic = (c2 & 32u) != 0;   // a resting partner is position only (rest_ball(), bcu_pipeline.cu)
                const float2 x1 = w.pos[g1], y1 = w.vel[g1], z1 = w.acc[g1], x2 = w.pos[g2];
                float2 y2 = make_float2(0.f, 0.f), z2 = make_float2(0.f, 0.f);
                if (partnerDynamic) { y2 = w.vel[g2]; z2 = w.acc[g2]; }
                const v2 a2 = partnerDynamic ? mk(z2.x, z2.y) : rest_acc(p_phys);
                float co[5];
                collision_coefficients(mk(z1.x, z1.y), mk(y1.x, y1.y), mk(x1.x, x1.y), a2, mk(y2.x, y2.y),
                                       mk(x2.x, x2.y), diameter, co);
                const uint64_t j = (uint64_t)sJobBase + iq;
                if (j < w.jobCap) {
                    #pragma unroll
                    for (int q = 0; q < 5; ++q) w.jobCo[(size_t)q * w.jobCap + j] = co[q];
                    w.jobCode[j] = (en << 6) | (c2 & 63u) | ((c2 & 0x8000u) ? JOB_MARK : 0u);
                } else {
                    w.counts[4 * ROUNDS + 3] = 1u;  // queue overflow: the host grows the queue and re-runs the launch
                }
            }
        }
#endif
        __syncthreads();
    }
    if (p.out.counters && skipped) atomicAdd(&p.out.counters[5], skipped);
}

__global__ void __launch_bounds__(ITEM_BLOCK, ITEM_MIN_BLOCKS) items_kernel(const LaunchParams* __restrict__ pp, PipeWorkspace w, int round) {
    extern __shared__ uint16_t smemQ[];
    __shared__ ItemShared sh;
    if (w.pdlEarly) cudaTriggerProgrammaticLaunchCompletion();
    items_prologue(sh, w);
    cudaGridDependencySynchronize();   // nothing above depends on the previous kernel
    __syncthreads();
    items_body(sh, smemQ, *pp, w, round, blockIdx.x, gridDim.x);
}
// The same kernel compiled for more resident CTAs per SM (fewer registers per thread): slower per thread, which is what a
// small, latency-bound round pays for, and faster in aggregate once a round has enough work to be issue-bound.
__global__ void __launch_bounds__(ITEM_BLOCK, ITEM_DENSE_MIN_BLOCKS) items_kernel_dense(const LaunchParams* __restrict__ pp, PipeWorkspace w, int round) {
    extern __shared__ uint16_t smemQ[];
    __shared__ ItemShared sh;
    if (w.pdlEarly) cudaTriggerProgrammaticLaunchCompletion();
    items_prologue(sh, w);
    cudaGridDependencySynchronize();
    __syncthreads();
    items_body(sh, smemQ, *pp, w, round, blockIdx.x, gridDim.x);
}
