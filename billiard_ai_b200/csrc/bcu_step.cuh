// step_kernel: the "fused" layout of a round (BCU_LAYOUT=1; included by bcu_pipeline.cu inside its anonymous namespace, after
// bcu_items.cuh whose reach bound, filters and constants it shares).  AN EXPERIMENT THAT IS KEPT, NOT THE DEFAULT: bit-exact
// (the GPU tests pass with it), 3-4 x slower than the three-kernel round.
//
// Idea: a GROUP of LANES lanes of one warp owns a STRIKE for a whole step,
//
//   step k = apply event k of the strike (createLayer :1957-1992, System::append type.cpp:467-531; k >= 1)
//            + next-event candidates of the new layer (nextRailCollision :1639-1682, nextPocketCollision :1684-1714,
//              nextBallCollision :1603-1637, nextBallInRest :1590, nextRolling :1575; k <= MAX_EVENTS)
//
// with the strike's balls staged in shared memory (loaded once per step) and the rails, pockets, resting partners and moving
// pairs of all its dynamic balls spread over the lanes of the group: a round is two kernels (step, quartic) instead of
// three, the resting balls are loaded once per strike instead of once per dynamic ball, the strike's earliest rest /
// rolling / rail / pocket time is a min over the group's registers and one plain store of the key.
//
// Measured (B200, 65 536 strikes, 22 balls, profiles/r02_step_kernel_peak_ncu_full.txt): 4.6 / 5.8 / 8.3 ms per level with 4 / 8 /
// 16 lanes per strike against 1.6 ms.  A step at the peak of a level executes 83 M warp-instructions with 7.8 of 32 lanes
// active (items + apply of the three-kernel round: 15.5 M at 26 lanes): the heavy second stages -- exact rail / pocket
// tests, moving-vs-moving pre-checks, coefficients, the event itself -- have a handful of items per STRIKE, so most lanes of
// a group idle through them, and every lane of a group repeats the per-ball scalar work (lengths, reach, square roots).  The
// three-kernel round compacts exactly those stages across a whole CTA (and the lists across the grid), which is what keeps
// its lanes busy; the kernel boundaries are the price.  What this experiment settles: the round's floor is not a per-thread
// instruction chain that lane-splitting could shorten.
#pragma once

#ifndef STEP_BLOCK_V
#define STEP_BLOCK_V 128
#endif
constexpr int STEP_BLOCK = STEP_BLOCK_V;
#ifndef STEP_MIN_BLOCKS
#define STEP_MIN_BLOCKS 5
#endif
constexpr int STEP_MAX_GROUPS = STEP_BLOCK / 4;

struct StepShared {
    uint32_t groupJobs[STEP_MAX_GROUPS];    // pairs pushed by each group in this tile
    uint32_t groupBase[STEP_MAX_GROUPS];    // first global job slot of the group's queue
    uint32_t groupLiveSlot[STEP_MAX_GROUPS];// slot in the next live list, ~0u: strike finished
    RailGeom rails[MAX_RAILS];
    float detQV[MAX_RAILS];
    float railLenSq[MAX_RAILS];
    PocketGeom pockets[MAX_POCKETS];
};

// shared-memory staging of one strike (a group's slice of the dynamic shared memory): per ball slot
struct GroupStage {
    float2* pos;        // every ball on the table (resting balls: only this and meta are loaded)
    float2* vel;
    float2* vn;
    float2* acc;
    float* nextRolling;
    uint32_t* meta;
    uint32_t* line;     // lineCand word of the ball (bcu_items.cuh: line cache)
    uint16_t* queue;    // pairs that need the quartic: ball << 6 | partner is dynamic << 5 | partner, bit 15 = audit mark
};
__host__ __device__ inline size_t step_group_bytes(int nBalls) {
    const size_t perBall = 4 * sizeof(float2) + sizeof(float) + 2 * sizeof(uint32_t);
    return ((size_t)nBalls * perBall + (size_t)2 * nBalls * sizeof(uint16_t) + 15) & ~(size_t)15;
}
__device__ __forceinline__ GroupStage step_group_stage(unsigned char* base, int nBalls) {
    GroupStage g;
    g.pos = reinterpret_cast<float2*>(base);
    g.vel = g.pos + nBalls; g.vn = g.vel + nBalls; g.acc = g.vn + nBalls;
    g.nextRolling = reinterpret_cast<float*>(g.acc + nBalls);
    g.meta = reinterpret_cast<uint32_t*>(g.nextRolling + nBalls);
    g.line = g.meta + nBalls;
    g.queue = reinterpret_cast<uint16_t*>(g.line + nBalls);
    return g;
}
__device__ __forceinline__ BallRec stage_load(const GroupStage& g, int b) {
    BallRec r;
    const float2 q0 = g.pos[b], q1 = g.vel[b], q2 = g.vn[b], q3 = g.acc[b];
    const uint32_t m = g.meta[b];
    r.pos = mk(q0.x, q0.y); r.vel = mk(q1.x, q1.y); r.vn = mk(q2.x, q2.y); r.acc = mk(q3.x, q3.y);
    r.nextRolling = g.nextRolling[b];
    r.kind = meta_kind(m); r.rolling = meta_rolling(m); r.lastRail = meta_last_rail(m); r.type = meta_type(m);
    return r;
}
// new state of a ball: to the stage (the items phase of this step reads it) and to the workspace (the next step / finish)
__device__ __forceinline__ void stage_store(const GroupStage& g, const PipeWorkspace& w, int b, int64_t s, const BallRec& r) {
    g.pos[b] = make_float2(r.pos.x, r.pos.y);
    g.vel[b] = make_float2(r.vel.x, r.vel.y);
    g.vn[b] = make_float2(r.vn.x, r.vn.y);
    g.acc[b] = make_float2(r.acc.x, r.acc.y);
    g.nextRolling[b] = r.nextRolling;
    g.meta[b] = make_meta(r.kind, r.rolling, r.lastRail, r.type);
    store_ball(w, (int64_t)b * w.n + s, r);
}
__device__ __forceinline__ void consider(unsigned long long& best, float t, uint32_t code) {
    if (t != t) return;   // a NaN time never wins the reference's strict '<'
    const unsigned long long k2 = ((unsigned long long)ordered_bits(t) << 32) | code;
    best = k2 < best ? k2 : best;
}
__device__ __forceinline__ void consider_own(unsigned long long& best, const BallRec& r, int b) {   // nextBallInRest :1590, nextRolling :1575
    consider(best, time_to_stop(r.acc, r.vel), make_code(b, CAT_REST, 0));
    if (r.nextRolling >= 0.0f) consider(best, r.nextRolling, make_code(b, CAT_ROLLING, 0));
}
template <int LANES>
__device__ __forceinline__ unsigned long long group_min(unsigned long long v, unsigned gmask) {
    #pragma unroll
    for (int d = LANES / 2; d >= 1; d >>= 1) {
        const unsigned long long o = __shfl_xor_sync(gmask, v, d);
        v = o < v ? o : v;
    }
    return v;
}
template <int LANES>
__device__ __forceinline__ uint32_t group_or(uint32_t v, unsigned gmask) {
    #pragma unroll
    for (int d = LANES / 2; d >= 1; d >>= 1) v |= __shfl_xor_sync(gmask, v, d);
    return v;
}

// quantised coefficients of one queued pair -> global job queue (stage 2 of bcu_items.cuh)
__device__ __forceinline__ void step_post_job(const GroupStage& g, const PipeWorkspace& w, int64_t s, uint32_t code16, uint64_t j,
                                              float diameter, v2 restAcc) {
    const int o = (int)(code16 & 31u), b = (int)((code16 >> 6) & 31u);
    const bool partnerDynamic = (code16 & 32u) != 0;
    const float2 x1 = g.pos[b], y1 = g.vel[b], z1 = g.acc[b], x2 = g.pos[o];
    const v2 v2o = partnerDynamic ? mk(g.vel[o].x, g.vel[o].y) : mk(0.0f, 0.0f);
    const v2 a2o = partnerDynamic ? mk(g.acc[o].x, g.acc[o].y) : restAcc;
    float co[5];
    collision_coefficients(mk(z1.x, z1.y), mk(y1.x, y1.y), mk(x1.x, x1.y), a2o, v2o, mk(x2.x, x2.y), diameter, co);
    if (j < w.jobCap) {
        #pragma unroll
        for (int q = 0; q < 5; ++q) w.jobCo[(size_t)q * w.jobCap + j] = co[q];
        w.jobCode[j] = ((((uint32_t)s << 5) | (uint32_t)b) << 6) | (code16 & 63u) | ((code16 & 0x8000u) ? JOB_MARK : 0u);
    } else {
        w.counts[4 * ROUNDS + 3] = 1u;  // queue overflow: the host grows the queue and re-runs the launch
    }
}

template <int LANES>
__device__ __forceinline__ void step_body(StepShared& sh, unsigned char* stageMem, const LaunchParams& p, const PipeWorkspace& w, int k,
                                          unsigned bid, unsigned nb) {
    constexpr int GROUPS = STEP_BLOCK / LANES;
    const PhysConst pc = p_phys;
    const int tid = threadIdx.x, lane = tid & 31, gl = tid & (LANES - 1), grp = tid / LANES;
    const unsigned gmask = LANES == 32 ? 0xffffffffu : (((1u << LANES) - 1u) << (lane & ~(LANES - 1)));
    const int R = p.table->nRails, P = p.table->nPockets, NB = p.states.nBalls;
    const uint32_t nLive = w.counts[4 * k + 0];
    if (nLive == 0) return;
    const int64_t n = w.n;
    const uint32_t* live = w.live + (size_t)(k & 1) * n;
    uint32_t* liveNext = w.live + (size_t)((k + 1) & 1) * n;
    uint32_t* liveCountNext = &w.counts[4 * (k + 1) + 0];
    uint32_t* jobCount = &w.counts[4 * k + 2];
    const bool doApply = k >= 1, doItems = k < ROUNDS;
    const int round = k - 1;                        // the event applied here is event `round + 1` = k
    const bool overflow = round + 1 > MAX_EVENTS;   // billiard_search.cpp:1376-1379
    const float diameter = p.table->diameter, dd = diameter * diameter;
    const int prune = w.pruneMode;
    const int QCAP = 2 * NB;
    const v2 restAcc = pc.slideAccLen * mk(0.0f, 0.0f);   // BallState of a ball at rest: acceleration = length * (0, 0) = (-0, -0)
    const GroupStage g = step_group_stage(stageMem + (size_t)grp * step_group_bytes(NB), NB);
    const RailGeom* sRails = sh.rails;
    const PocketGeom* sPockets = sh.pockets;
    unsigned long long skipped = 0;   // counters / audit only

    for (uint32_t base = bid * GROUPS; base < nLive; base += nb * GROUPS) {
        if (tid < GROUPS) { sh.groupJobs[tid] = 0; sh.groupLiveSlot[tid] = ~0u; }
        __syncthreads();
        const uint32_t idx = base + grp;
        bool alive = false;
        int64_t s = 0;
        if (idx < nLive) {
            s = live[idx];
            const unsigned long long key = w.key[s];
            float T = w.T[s];
            uint32_t dynMask = w.dynMask[s], restMask = w.restMask[s];
            // ---- stage the strike: every ball's position and meta word, the full state of the dynamic ones (one round trip) ----
            for (int b = gl; b < NB; b += LANES) {
                const uint32_t bit = 1u << b;
                if (!((dynMask | restMask) & bit)) continue;
                const int64_t gi = (int64_t)b * n + s;
                g.pos[b] = w.pos[gi];
                g.meta[b] = w.meta[gi];
                if (dynMask & bit) {
                    g.vel[b] = w.vel[gi]; g.vn[b] = w.vn[gi]; g.acc[b] = w.acc[gi];
                    g.nextRolling[b] = w.nextRolling[gi];
                    g.line[b] = w.lineCand[gi];
                }
            }
            __syncwarp(gmask);
            unsigned long long best = ~0ull;
            alive = true;
            if (!doApply) {
                if (gl == 0) best = key;   // step 0: the rest / rolling offers init_kernel made for layer 0
            } else if (key == ~0ull) {
                if (gl == 0) mark_finished(p, s, F_NOEVENT, round, T);              // billiard_search.cpp:1464-1467
                alive = false;
            } else if (overflow) {
                if (gl == 0) mark_finished(p, s, F_OVERFLOW, round, T);
                alive = false;
            } else {
                // ---- the event of the strike (createLayer :1957-1992) ------------------------------------------------------
                const float dt = from_ordered_bits((uint32_t)(key >> 32));
                const uint32_t code = (uint32_t)key;
                const int evBall = (int)(code >> 16);
                const uint32_t cat = (code >> 8) & 0xffu;
                const int evPartner = (int)(code & 0xffu);
                const int evType = cat == CAT_REST ? E_REST : cat == CAT_ROLLING ? E_ROLLING : cat == CAT_RAIL ? E_RAIL_COLLISION
                                 : cat == CAT_POCKET ? E_POTTING : E_BALL_COLLISION;
                const bool pairEv = cat == CAT_STATIC_PAIR || cat == CAT_DYNAMIC_PAIR;
                const float newT = T + dt;
                const float deltaTime = newT - T;  // type.cpp:475
                const uint32_t others = dynMask & ~(1u << evBall) & ~(pairEv ? (1u << evPartner) : 0u);
                if (gl == 0) {
                    BallRec A = stage_load(g, evBall);
                    const v2 npos = advance_position(A.acc, A.vel, dt, A.pos);
                    const v2 nvel = advance_velocity(A.acc, A.vel, dt);
                    int cause = -1;
                    A.pos = npos;
                    if (evType == E_POTTING) {                 // :1716-1746
                        A.vel = mk(0.f, 0.f); A.rolling = false; A.kind = K_POTTING;
                    } else if (evType == E_RAIL_COLLISION) {   // :1748-1786
                        const RailGeom& rg = sRails[evPartner];
                        A.vel = rail_collision(nvel, mk(rg.nx, rg.ny), pc);
                        A.rolling = true; A.kind = K_RAIL; cause = evPartner;
                    } else if (evType == E_REST) {             // :1788-1808
                        A.vel = mk(0.f, 0.f); A.rolling = false; A.kind = K_REST;
                    } else if (evType == E_ROLLING) {          // :1810-1840
                        A.vel = nvel; A.rolling = true; A.kind = K_MOVING;
                    } else {                                   // :1842-1905 createBallCollisionNodes
                        BallRec B;
                        if (cat == CAT_DYNAMIC_PAIR) {
                            B = stage_load(g, evPartner);
                        } else {   // a ball at rest: zero velocity, (-0, -0) acceleration, no rolling time (only position and meta are staged)
                            const float2 q = g.pos[evPartner];
                            const uint32_t m = g.meta[evPartner];
                            B.pos = mk(q.x, q.y); B.vel = mk(0.f, 0.f); B.vn = mk(0.f, 0.f); B.acc = restAcc; B.nextRolling = -1.0f;
                            B.kind = meta_kind(m); B.rolling = false; B.lastRail = meta_last_rail(m); B.type = meta_type(m);
                        }
                        const v2 opos = advance_position(B.acc, B.vel, dt, B.pos);
                        const v2 ovel = advance_velocity(B.acc, B.vel, dt);
                        v2 o1, o2;
                        elastic_collision(npos, nvel, opos, ovel, A.rolling, B.rolling, pc, &o1, &o2);
                        A.vel = o1; A.kind = K_COLLISION;
                        B.pos = opos; B.vel = o2; B.kind = K_COLLISION;
                        derive_ball(B.vel, B.rolling, pc, &B.vn, &B.acc);
                        append_bookkeeping(B, deltaTime, -1, pc);
                        stage_store(g, w, evPartner, s, B);
                        g.line[evPartner] = 0u;
                        consider_own(best, B, evPartner);
                    }
                    derive_ball(A.vel, A.rolling, pc, &A.vn, &A.acc);
                    append_bookkeeping(A, deltaTime, cause, pc);
                    stage_store(g, w, evBall, s, A);
                    if (kind_dynamic(A.kind)) consider_own(best, A, evBall);
                    if (p.out.eventLog) {
                        int a = evBall, b2 = evPartner;
                        if (evType == E_BALL_COLLISION && a > b2) { const int t = a; a = b2; b2 = t; }
                        if (evType == E_REST || evType == E_ROLLING) b2 = -1;
                        p.out.eventLog[s * 32 + round] = make_int4(evType, a, b2, __float_as_int(newT));
                    }
                }
                // ---- every other dynamic ball: closed-form advance to the new layer (createNodes :1907-1955) ---------------
                {
                    int ord = 0;
                    for (uint32_t m = others; m; m &= m - 1, ++ord) {
                        if ((ord & (LANES - 1)) != gl) continue;
                        const int b = __ffs(m) - 1;
                        BallRec r = stage_load(g, b);
                        const v2 nv = advance_velocity(r.acc, r.vel, dt);
                        r.pos = advance_position(r.acc, r.vel, dt, r.pos);
                        r.vel = nv; r.kind = K_MOVING;
                        derive_ball(r.vel, r.rolling, pc, &r.vn, &r.acc);
                        append_bookkeeping(r, deltaTime, -1, pc);
                        stage_store(g, w, b, s, r);
                        consider_own(best, r, b);
                    }
                }
                // the masks of the new layer follow from the event alone (every lane of the group computes them)
                const bool stays = evType == E_RAIL_COLLISION || evType == E_ROLLING || evType == E_BALL_COLLISION;
                if (pairEv) { dynMask |= 1u << evPartner; restMask &= ~(1u << evPartner); }
                if (!stays) {
                    dynMask &= ~(1u << evBall);
                    if (evType == E_REST) restMask |= 1u << evBall;
                }
                T = newT;
                if (gl == 0) {
                    w.dynMask[s] = dynMask; w.restMask[s] = restMask; w.T[s] = newT;
                    if (dynMask == 0) mark_finished(p, s, F_COMPLETED, round + 1, newT);
                }
                if (dynMask == 0) alive = false;
                __syncwarp(gmask);
            }

            if (alive && !doItems) alive = false;   // (step ROUNDS only marks: every strike still live has overflowed or has no event)
            if (alive) {
                if (p.out.counters && gl == 0) atomicAdd(&w.counts[4 * k + 1], (uint32_t)__popc(dynMask));
                // ---- rails and pockets of every dynamic ball, spread over the lanes of the group ---------------------------
                for (uint32_t m = dynMask; m; m &= m - 1) {
                    const int b = __ffs(m) - 1;
                    const float2 q0 = g.pos[b], q1 = g.vel[b], q2n = g.vn[b], q3 = g.acc[b];
                    const uint32_t metaWord = g.meta[b];
                    const v2 pos = mk(q0.x, q0.y), vel = mk(q1.x, q1.y), dn = mk(q2n.x, q2n.y), acc = mk(q3.x, q3.y);
                    if (is_zero(vel)) continue;
                    const int lastRail = meta_last_rail(metaWord);
                    uint32_t railMask = 0, pocketMask = 0;
                    // line cache (bcu_items.cuh): a ball that was merely advanced reuses the masks of its line of motion
                    const bool reuse = prune && meta_kind(metaWord) == K_MOVING && R <= 24;
                    if (reuse) {
                        const uint32_t c2 = g.line[b];
                        railMask = c2 & 0xffffffu;
                        pocketMask = c2 >> 24;
                    } else {
                        const float speedSq = dot(vel, vel);
                        for (int r = gl; r < R; r += LANES) {   // pass A (no division), see bcu_items.cuh
                            const RailGeom rg = sRails[r];
                            const float detDV = vel.x * rg.dy - vel.y * rg.dx;
                            const float nT = (rg.dx * pos.y - rg.dy * pos.x) + sh.detQV[r];
                            bool cand = r != lastRail && detDV != 0.0f;
                            if (prune) {
                                const bool behind = nT * detDV < 0.0f && fabsf(nT) > 1e-10f * fabsf(detDV);
                                const float lenSq = sh.railLenSq[r];
                                const float nU = (pos.x - rg.sx) * vel.y - (pos.y - rg.sy) * vel.x;   // u = nU / -detDV
                                const float den = -detDV, sgnU = den < 0.0f ? -nU : nU, aden = fabsf(den);
                                const bool wellCond = lenSq > 0.0f && detDV * detDV >= 4e-4f * (speedSq * lenSq);
                                const bool outside = wellCond && (sgnU < -0.05f * aden || sgnU > 1.05f * aden);
                                cand = cand && !behind && !outside;
                            }
                            if (cand) railMask |= 1u << r;
                        }
                        for (int kk = gl; kk < P; kk += LANES) {
                            const PocketGeom pg = sPockets[kk];
                            const v2 delta = pos - mk(pg.x, pg.y);
                            const float qa = 1;
                            const float qb = 2 * dot(dn, delta);
                            const float qc = dot(delta, delta) - pg.r * pg.r;
                            const float disc = qb * qb - 4 * qa * qc;
                            const float slack = prune ? 1e-5f * (qb * qb + 4.0f * (dot(delta, delta) + pg.r * pg.r)) + 1.0f : 0.0f;
                            if (!(disc < -slack) && !(prune && qb > 0.0f && qc > 0.0f)) pocketMask |= 1u << kk;
                        }
                        railMask = group_or<LANES>(railMask, gmask);
                        pocketMask = group_or<LANES>(pocketMask, gmask);
                        if (prune && R <= 24 && gl == 0) w.lineCand[(int64_t)b * n + s] = railMask | (pocketMask << 24);
                    }
                    const float len1 = length(vel);
                    const float accLen = length(acc);
                    // exact stages of the candidates (timeToCollisionWithRail :181-238, timeToTarget :383-457), one per lane
                    int ord = 0;
                    for (uint32_t c = railMask; c; c &= c - 1, ++ord) {
                        if ((ord & (LANES - 1)) != gl) continue;
                        const int r = __ffs(c) - 1;
                        const RailGeom rg = sRails[r];
                        const float detDV = vel.x * rg.dy - vel.y * rg.dx;
                        const float detVP = rg.dx * pos.y - rg.dy * pos.x;
                        const float lambda = (detVP + sh.detQV[r]) / detDV;
                        const bool useX = rg.dx != 0.0f;
                        const float num = useX ? (pos.x + lambda * vel.x - rg.sx) : (pos.y + lambda * vel.y - rg.sy);
                        const float u = num / (useX ? rg.dx : rg.dy);
                        if (lambda < 0.0f || u < 0.0f || u > 1.0f) continue;
                        const float distance = length(lambda * vel);
                        float x[2];
                        if (solve_quadratic(0.5f * -accLen, len1, -distance, x) != 0) consider(best, x[0], make_code(b, CAT_RAIL, r));
                    }
                    for (uint32_t c = pocketMask; c; c &= c - 1, ++ord) {
                        if ((ord & (LANES - 1)) != gl) continue;
                        const int kk = __ffs(c) - 1;
                        if (reuse) {   // passed since the mask was made? (outside and moving away: none)
                            const PocketGeom pg = sPockets[kk];
                            const v2 delta = pos - mk(pg.x, pg.y);
                            if (2 * dot(dn, delta) > 0.0f && dot(delta, delta) - pg.r * pg.r > 0.0f) continue;
                        }
                        float t;
                        if (time_to_pocket(sPockets[kk], pos, dn, vel, acc, &t)) consider(best, t, make_code(b, CAT_POCKET, kk));
                    }
                }
                // the strike's earliest rest / rolling / rail / pocket event: its key for this round, and the T of the reach bound
                best = group_min<LANES>(best, gmask);
                if (gl == 0) w.key[s] = best;
                const float tBest = from_ordered_bits((uint32_t)(best >> 32));   // NaN when the strike has no offer at all
                const Reach rc = reach_of(tBest, diameter);
                const float corridor = diameter + 0.5f;
                // ---- resting partners of every dynamic ball (:318-324) --------------------------------------------------------
                for (uint32_t m = dynMask; m; m &= m - 1) {
                    const int b = __ffs(m) - 1;
                    const float2 q0 = g.pos[b], q1 = g.vel[b], q2n = g.vn[b], q3 = g.acc[b];
                    const v2 pos = mk(q0.x, q0.y), vel = mk(q1.x, q1.y), dn = mk(q2n.x, q2n.y), acc = mk(q3.x, q3.y);
                    if (is_zero(vel)) continue;
                    const float len1 = length(vel), accLen = length(acc);
                    const float reach = rc.fixed + reach_ball(rc, len1, accLen);
                    const float reachSq = reach * reach;
                    const float passMax = 0.99f * (len1 * len1) / (2.0f * accLen) - diameter - 5.0f;
                    const float passMaxSq = passMax > 0.0f ? passMax * passMax : -1.0f;
                    for (int o = gl; o < NB; o += LANES) {
                        if (!((restMask >> o) & 1u)) continue;
                        const float2 q2 = g.pos[o];
                        if (prune) {   // pass A: farther than d + 0.5 mm from the line of motion fails the corridor test whichever arm it takes
                            const float cr = dn.x * (q2.y - pos.y) - dn.y * (q2.x - pos.x);
                            if (fabsf(cr) > corridor) continue;
                        }
                        const v2 toPoint = mk(q2.x, q2.y) - pos;
                        const float factor = dot(dn, toPoint) / len1;
                        const v2 dv = toPoint - (vel * factor);
                        const float distSq = dot(toPoint, toPoint);
                        const float dist = factor < 0 ? distSq : dot(dv, dv);
                        if (dist > dd) continue;
                        bool skip = false;
                        if (prune) {
#if PRUNE_REST_PASS_ONLY_V
                            if (distSq <= passMaxSq) skip = distSq > reachSq;   // passed completely before the turning point
#else
                            skip = distSq > reachSq;   // guard variant for tests/golden/prune_regressions.json: NOT sound
#endif
                        }
                        uint32_t code16 = ((uint32_t)b << 6) | (uint32_t)o;
                        if (skip) {
                            ++skipped;
                            if (prune != 2) continue;
                            code16 |= 0x8000u;
                        }
                        const uint32_t at = atomicAdd(&sh.groupJobs[grp], 1u);
                        if (at < (uint32_t)QCAP) g.queue[at] = (uint16_t)code16;
                        else step_post_job(g, w, s, code16, (uint64_t)atomicAdd(jobCount, 1u), diameter, restAcc);
                    }
                }
                // ---- moving-vs-moving pairs (:292-317), one job covers both orders -----------------------------------------
                {
                    int ord = 0;
                    for (uint32_t m = dynMask; m; m &= m - 1) {
                        const int b = __ffs(m) - 1;
                        for (uint32_t pm = dynMask & ~((2u << b) - 1u); pm; pm &= pm - 1, ++ord) {
                            if ((ord & (LANES - 1)) != gl) continue;
                            const int o = __ffs(pm) - 1;
                            const float2 x1 = g.pos[b], y1 = g.vel[b], n1 = g.vn[b], z1 = g.acc[b];
                            const float2 x2 = g.pos[o], y2 = g.vel[o], n2 = g.vn[o], z2 = g.acc[o];
                            const v2 p1 = mk(x1.x, x1.y), v1 = mk(y1.x, y1.y), p2 = mk(x2.x, x2.y), w2 = mk(y2.x, y2.y);
                            const v2 v1n = mk(n1.x, n1.y), w2n = mk(n2.x, n2.y);
                            bool outOfReach = false;
                            if (prune) {
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
                            bool pass = false;
                            if (!outOfReach || prune == 2) {
                                #pragma unroll 1
                                for (int turn = 0; turn < 2 && !pass; ++turn) {   // the reference tests b -> o and, on o's own turn, o -> b
                                    const v2 va = turn ? w2 : v1, pa = turn ? p2 : p1, van = turn ? w2n : v1n;
                                    const v2 vb = turn ? v1 : w2, pb = turn ? p1 : p2, vbn = turn ? v1n : w2n;
                                    if (is_zero(va)) continue;                                  // :287
                                    if (dot(vb, vb) > 0) pass = collision_precheck_moving(va, pa, van, vb, pb, vbn, diameter);
                                    else pass = !(point_to_half_line_sq(pa, va, pb) > diameter * diameter);   // :318-324
                                }
                            }
                            uint32_t code16 = ((uint32_t)b << 6) | 32u | (uint32_t)o;
                            if (pass && outOfReach) { ++skipped; code16 |= 0x8000u; }   // audit only
                            else if (outOfReach && prune != 2) ++skipped;
                            if (!pass) continue;
                            const uint32_t at = atomicAdd(&sh.groupJobs[grp], 1u);
                            if (at < (uint32_t)QCAP) g.queue[at] = (uint16_t)code16;
                            else step_post_job(g, w, s, code16, (uint64_t)atomicAdd(jobCount, 1u), diameter, restAcc);
                        }
                    }
                }
                if (gl == 0) sh.groupLiveSlot[grp] = 0u;   // still live: gets its slot below
            }
        }
        __syncthreads();
        // ---- one reservation per tile: slots of the next live list and of the global job queue ------------------------------
        if (tid == 0) {
            uint32_t jobs = 0, lives = 0;
            for (int i = 0; i < GROUPS; ++i) {
                const uint32_t c = sh.groupJobs[i] < (uint32_t)QCAP ? sh.groupJobs[i] : (uint32_t)QCAP;
                sh.groupJobs[i] = c;
                sh.groupBase[i] = jobs; jobs += c;
                if (sh.groupLiveSlot[i] != ~0u) sh.groupLiveSlot[i] = lives++;
            }
            const uint32_t jb = jobs ? atomicAdd(jobCount, jobs) : 0u;
            const uint32_t lb = lives ? atomicAdd(liveCountNext, lives) : 0u;
            for (int i = 0; i < GROUPS; ++i) {
                sh.groupBase[i] += jb;
                if (sh.groupLiveSlot[i] != ~0u) sh.groupLiveSlot[i] += lb;
            }
        }
        __syncthreads();
        if (alive) {
            if (gl == 0) liveNext[sh.groupLiveSlot[grp]] = (uint32_t)s;
            const uint32_t cnt = sh.groupJobs[grp], jb = sh.groupBase[grp];
            for (uint32_t i = gl; i < cnt; i += LANES) step_post_job(g, w, s, g.queue[i], (uint64_t)jb + i, diameter, restAcc);
        }
        __syncthreads();
    }
    if (p.out.counters && skipped) atomicAdd(&p.out.counters[5], skipped);
}

template <int LANES>
__global__ void __launch_bounds__(STEP_BLOCK, STEP_MIN_BLOCKS) step_kernel(const LaunchParams* __restrict__ pp, PipeWorkspace w, int k) {
    extern __shared__ __align__(16) unsigned char stageMem[];
    __shared__ StepShared sh;
    if (w.pdlEarly) cudaTriggerProgrammaticLaunchCompletion();
    {   // table geometry -> shared memory; depends on nothing a kernel of the launch writes
        const LaunchParams& p = *pp;
        const int tid = threadIdx.x;
        const int R = p.table->nRails, P = p.table->nPockets;
        for (int i = tid; i < R; i += STEP_BLOCK) {
            const RailGeom rg = p.table->rails[i];
            sh.rails[i] = rg;
            sh.detQV[i] = rg.sx * rg.dy - rg.sy * rg.dx;
            const float used = fabsf(rg.dx != 0.0f ? rg.dx : rg.dy);
            sh.railLenSq[i] = used >= 5.0f ? rg.dx * rg.dx + rg.dy * rg.dy : -1.0f;
        }
        for (int i = tid; i < P; i += STEP_BLOCK) sh.pockets[i] = p.table->pockets[i];
    }
    cudaGridDependencySynchronize();
    __syncthreads();
    step_body<LANES>(sh, stageMem, *pp, w, k, blockIdx.x, gridDim.x);
}
