// round_kernel: the "merged" layout of a round (BCU_LAYOUT=2; included by bcu_pipeline.cu inside its anonymous namespace, after
// bcu_items.cuh whose reach bound, filters, queues and constants it shares).  AN EXPERIMENT THAT IS KEPT, NOT THE DEFAULT:
// bit-exact (the GPU tests pass with it), 30 % slower than the three-kernel round.
//
// Idea: apply_kernel is the cheapest of the three kernels of a round to give up -- everything it computes is a pure function of
// the PREVIOUS layer and the winning event, and the thread that owns a dynamic ball in items_kernel loads that ball anyway.  So
//
//   round k = [ apply event k of every live strike ] + [ next-event candidates of layer k ]      (k = 0: candidates only)
//
// is ONE kernel with the thread-per-dynamic-ball mapping of items_kernel: the state lives in two buffers (layer k - 1 is
// read-only while layer k is written), and every thread derives what it needs of layer k from layer k - 1 by itself --
// its own ball (createLayer :1957-1992 for the event ball(s), the closed-form advance :1907-1955 for the others), and, in
// the pair stages, the partner's (the same arithmetic again).  The strike's bookkeeping (masks, layer time, event log,
// finish marks) is done by the thread of the event ball; a resting ball that is hit becomes an extra item of the tile, taken
// by a spare warp; a ball that stops is carried over to the other buffer one round later.  A round is two kernels.
//
// Measured (B200, 22 balls): 28.2 M strikes/s against 41.0 M on the 65 536-strike level, and -- the number that matters for
// the idea -- a 256-strike level takes 1.04 ms against 0.92 ms.  One kernel fewer per round made the fixed cost of the chain
// LARGER: what a small round waits for is the sum of its dependent memory round trips, barriers and arithmetic chains, and
// the merged kernel's chain (list entry -> key, layer time and masks -> the event's balls -> ... and the event decoded again
// in both pair stages) is longer than items' and apply's put together, while a kernel boundary under a CUDA graph with
// programmatic dependent launch costs almost nothing.  With the fused step kernel (bcu_step.cuh) this is the second layout
// that removes kernel boundaries and loses; the three-kernel round stays.
#pragma once

constexpr int RK_BLOCK = ITEM_BLOCK;          // 128
constexpr int RK_ENTRIES = RK_BLOCK - 32;     // dynamic-ball entries per tile; the last warp takes the tile's extra items
constexpr int RK_SLOTS = 2 * RK_ENTRIES;      // entries + at most one extra per entry
#ifndef RK_MIN_BLOCKS
#define RK_MIN_BLOCKS 5
#endif

struct StrikeEvent {
    bool has;          // an event is applied (k >= 1, key valid, no overflow)
    float dt, deltaTime, newT;
    int ball, partner, type;
    uint32_t cat;
    bool pair;
    uint32_t dynNew, restNew;
};
__device__ __forceinline__ StrikeEvent decode_event(unsigned long long key, float T, uint32_t dynOld, uint32_t restOld) {
    StrikeEvent ev;
    ev.has = true;
    ev.dt = from_ordered_bits((uint32_t)(key >> 32));
    const uint32_t code = (uint32_t)key;
    ev.ball = (int)(code >> 16);
    ev.cat = (code >> 8) & 0xffu;
    ev.partner = (int)(code & 0xffu);
    ev.type = ev.cat == CAT_REST ? E_REST : ev.cat == CAT_ROLLING ? E_ROLLING : ev.cat == CAT_RAIL ? E_RAIL_COLLISION
            : ev.cat == CAT_POCKET ? E_POTTING : E_BALL_COLLISION;
    ev.pair = ev.cat == CAT_STATIC_PAIR || ev.cat == CAT_DYNAMIC_PAIR;
    ev.newT = T + ev.dt;
    ev.deltaTime = ev.newT - T;   // type.cpp:475
    ev.dynNew = dynOld; ev.restNew = restOld;
    if (ev.pair) { ev.dynNew |= 1u << ev.partner; ev.restNew &= ~(1u << ev.partner); }
    const bool stays = ev.type == E_RAIL_COLLISION || ev.type == E_ROLLING || ev.type == E_BALL_COLLISION;
    if (!stays) {
        ev.dynNew &= ~(1u << ev.ball);
        if (ev.type == E_REST) ev.restNew |= 1u << ev.ball;
    }
    return ev;
}
__device__ __forceinline__ StrikeEvent no_event(uint32_t dynOld, uint32_t restOld) {
    StrikeEvent ev{};
    ev.has = false; ev.ball = -1; ev.partner = -1; ev.pair = false; ev.type = E_NONE;
    ev.dynNew = dynOld; ev.restNew = restOld;
    return ev;
}

__device__ __forceinline__ BallRec load_ball_from(const StateBuf& sb, int64_t g) {
    BallRec r;
    const float2 q0 = sb.pos[g], q1 = sb.vel[g], q2 = sb.vn[g], q3 = sb.acc[g];
    const uint32_t m = sb.meta[g];
    r.pos = mk(q0.x, q0.y); r.vel = mk(q1.x, q1.y); r.vn = mk(q2.x, q2.y); r.acc = mk(q3.x, q3.y);
    r.nextRolling = sb.nextRolling[g];
    r.kind = meta_kind(m); r.rolling = meta_rolling(m); r.lastRail = meta_last_rail(m); r.type = meta_type(m);
    return r;
}
__device__ __forceinline__ BallRec load_rest_ball_from(const StateBuf& sb, int64_t g, const PhysConst& pc) {
    BallRec r;
    const float2 q0 = sb.pos[g];
    const uint32_t m = sb.meta[g];
    r.pos = mk(q0.x, q0.y); r.vel = mk(0.f, 0.f); r.vn = mk(0.f, 0.f); r.acc = rest_acc(pc);
    r.nextRolling = -1.0f;
    r.kind = meta_kind(m); r.rolling = false; r.lastRail = meta_last_rail(m); r.type = meta_type(m);
    return r;
}
__device__ __forceinline__ void store_ball_to(const StateBuf& sb, int64_t g, const BallRec& r) {
    sb.pos[g] = make_float2(r.pos.x, r.pos.y);
    sb.vel[g] = make_float2(r.vel.x, r.vel.y);
    sb.vn[g] = make_float2(r.vn.x, r.vn.y);
    sb.acc[g] = make_float2(r.acc.x, r.acc.y);
    sb.nextRolling[g] = r.nextRolling;
    sb.meta[g] = make_meta(r.kind, r.rolling, r.lastRail, r.type);
}

// Ball x of strike s in layer k, derived from layer k - 1 (`old`) and the strike's event: exactly what apply_kernel writes
// for it.  x must be dynamic in layer k - 1 or the resting partner of a pair event.
__device__ __forceinline__ BallRec new_state(const StateBuf& old, int64_t n, int64_t s, int x, const StrikeEvent& ev, const PhysConst& pc,
                                             const RailGeom* sRails) {
    if (!ev.has) return load_ball_from(old, (int64_t)x * n + s);
    if (x == ev.ball || (ev.pair && x == ev.partner)) {
        BallRec A = load_ball_from(old, (int64_t)ev.ball * n + s);
        const v2 npos = advance_position(A.acc, A.vel, ev.dt, A.pos);
        const v2 nvel = advance_velocity(A.acc, A.vel, ev.dt);
        int cause = -1;
        A.pos = npos;
        if (ev.type == E_POTTING) {                 // :1716-1746
            A.vel = mk(0.f, 0.f); A.rolling = false; A.kind = K_POTTING;
        } else if (ev.type == E_RAIL_COLLISION) {   // :1748-1786
            const RailGeom& rg = sRails[ev.partner];
            A.vel = rail_collision(nvel, mk(rg.nx, rg.ny), pc);
            A.rolling = true; A.kind = K_RAIL; cause = ev.partner;
        } else if (ev.type == E_REST) {             // :1788-1808
            A.vel = mk(0.f, 0.f); A.rolling = false; A.kind = K_REST;
        } else if (ev.type == E_ROLLING) {          // :1810-1840
            A.vel = nvel; A.rolling = true; A.kind = K_MOVING;
        } else {                                    // :1842-1905 createBallCollisionNodes
            const int64_t gB = (int64_t)ev.partner * n + s;
            BallRec B = ev.cat == CAT_STATIC_PAIR ? load_rest_ball_from(old, gB, pc) : load_ball_from(old, gB);
            const v2 opos = advance_position(B.acc, B.vel, ev.dt, B.pos);
            const v2 ovel = advance_velocity(B.acc, B.vel, ev.dt);
            v2 o1, o2;
            elastic_collision(npos, nvel, opos, ovel, A.rolling, B.rolling, pc, &o1, &o2);
            A.vel = o1; A.kind = K_COLLISION;
            if (x == ev.partner) {
                B.pos = opos; B.vel = o2; B.kind = K_COLLISION;
                derive_ball(B.vel, B.rolling, pc, &B.vn, &B.acc);
                append_bookkeeping(B, ev.deltaTime, -1, pc);
                return B;
            }
        }
        derive_ball(A.vel, A.rolling, pc, &A.vn, &A.acc);
        append_bookkeeping(A, ev.deltaTime, cause, pc);
        return A;
    }
    BallRec r = load_ball_from(old, (int64_t)x * n + s);
    const v2 nv = advance_velocity(r.acc, r.vel, ev.dt);
    r.pos = advance_position(r.acc, r.vel, ev.dt, r.pos);
    r.vel = nv; r.kind = K_MOVING;
    derive_ball(r.vel, r.rolling, pc, &r.vn, &r.acc);
    append_bookkeeping(r, ev.deltaTime, -1, pc);
    return r;
}
// position of resting ball o of strike s in layer k (the event ball has just come to rest: its position is the advanced one)
__device__ __forceinline__ float2 rest_position(const StateBuf& old, int64_t n, int64_t s, int o, const StrikeEvent& ev) {
    const int64_t g = (int64_t)o * n + s;
    if (ev.has && ev.type == E_REST && o == ev.ball) {
        const float2 q0 = old.pos[g], q1 = old.vel[g], q3 = old.acc[g];
        const v2 np = advance_position(mk(q3.x, q3.y), mk(q1.x, q1.y), ev.dt, mk(q0.x, q0.y));
        return make_float2(np.x, np.y);
    }
    return old.pos[g];
}

struct RoundShared {
    ItemCounts cnt;
    uint32_t jobBase, nExtra, nDynAppend, nLiveAppend, dynBase, liveBase;
    uint32_t ent[RK_SLOTS];           // strike << 5 | ball of the item in this slot
    float T[RK_SLOTS];                // the strike's best event time so far when the item was processed (reach bound)
    uint32_t extra[RK_ENTRIES];       // resting balls hit in this tile's events: strike << 5 | ball
    uint32_t appendDyn[RK_SLOTS];     // items still dynamic in layer k: next round's list
    uint32_t appendLive[RK_ENTRIES];  // strikes still live
    float2 pos[RK_SLOTS], vel[RK_SLOTS], vn[RK_SLOTS], acc[RK_SLOTS];   // layer-k state of the item's ball (pair stages)
    RailGeom rails[MAX_RAILS];
    float detQV[MAX_RAILS];
    float railLenSq[MAX_RAILS];
    PocketGeom pockets[MAX_POCKETS];
};

// the strike's event as seen from layer k - 1 (memory: key, layer time, masks of the previous buffers)
__device__ __forceinline__ StrikeEvent strike_event(const PipeWorkspace& w, int k, int64_t s, bool* finishedOut, unsigned long long* keyOut,
                                                    float* TOut) {
    const int64_t n = w.n;
    if (k == 0) {
        *finishedOut = false; *keyOut = ~0ull; *TOut = w.T[s];
        return no_event(w.maskDyn[0][s], w.maskRest[0][s]);
    }
    const int prev = (k - 1) & 1;
    const unsigned long long key = (w.key + (size_t)((k - 1) % 3) * n)[s];
    const float T = (w.T + (size_t)prev * n)[s];
    const uint32_t dynOld = w.maskDyn[prev][s], restOld = w.maskRest[prev][s];
    *keyOut = key; *TOut = T;
    if (key == ~0ull || k > MAX_EVENTS) { *finishedOut = true; return no_event(dynOld, restOld); }
    *finishedOut = false;
    return decode_event(key, T, dynOld, restOld);
}

__device__ __forceinline__ void round_body(RoundShared& sh, uint16_t* smemQ, const LaunchParams& p, const PipeWorkspace& w, int k, unsigned bid,
                                           unsigned nb) {
    ItemCounts& cnt = sh.cnt;
    const RailGeom* sRails = sh.rails;
    const float* sDetQV = sh.detQV;
    const float* sRailLenSq = sh.railLenSq;
    const PocketGeom* sPockets = sh.pockets;
    const unsigned FULL = 0xffffffffu;
    const PhysConst pc = p_phys;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int R = w.nRails, P = w.nPockets, NB = w.nBalls;
    const int listIdx = k == 0 ? 0 : k - 1;
    const uint32_t nDyn = w.counts[4 * listIdx + 1];
    if (nDyn == 0) return;
    const int64_t n = w.n;
    const bool doApply = k >= 1, doItems = k < ROUNDS;
    const int prev = doApply ? (k - 1) & 1 : 0, cur = k & 1;
    const StateBuf old = w.buf[prev], neu = w.buf[cur];
    const uint32_t* dyn = w.dyn + (size_t)(listIdx & 1) * w.dynCap;
    uint32_t* dynNext = w.dyn + (size_t)(k & 1) * w.dynCap;
    uint32_t* liveNext = w.live + (size_t)(k & 1) * n;
    unsigned long long* keyCur = w.key + (size_t)(k % 3) * n;
    unsigned long long* keyStale = w.key + (size_t)((k + 1) % 3) * n;
    float* Tcur = w.T + (size_t)cur * n;
    uint32_t* jobCount = &w.counts[4 * k + 2];
    uint16_t* qPairs = smemQ;
    uint16_t* qDyn = qPairs + RK_SLOTS * NB;
    const float diameter = w.diameter;
    const float dd = diameter * diameter;
    const int prune = w.pruneMode;
    unsigned long long skipped = 0;   // counters / audit only

    for (uint32_t base = bid * RK_ENTRIES; base < nDyn; base += nb * RK_ENTRIES) {
        if (tid == 0) { cnt.pairs = 0; cnt.dynPairs = 0; sh.nExtra = 0; sh.nDynAppend = 0; sh.nLiveAppend = 0; }
        __syncthreads();
        // ---- phase 0: the entries of the tile; a resting ball that is hit in an entry's event becomes an extra item ----------
        uint32_t myEnt = 0;
        bool haveEntry = false;
        if (tid < RK_ENTRIES && base + tid < nDyn) {
            myEnt = dyn[base + tid];
            haveEntry = true;
            if (doApply) {
                const int64_t s = myEnt >> 5;
                const int b = (int)(myEnt & 31u);
                const unsigned long long key = (w.key + (size_t)((k - 1) % 3) * n)[s];
                if (key != ~0ull && k <= MAX_EVENTS) {
                    const uint32_t code = (uint32_t)key;
                    if ((int)(code >> 16) == b && ((code >> 8) & 0xffu) == CAT_STATIC_PAIR)
                        sh.extra[atomicAdd(&sh.nExtra, 1u)] = ((uint32_t)s << 5) | (code & 0xffu);
                }
            }
        }
        __syncthreads();
        const uint32_t nExtra = sh.nExtra;
        const int iters = warp < RK_ENTRIES / 32 ? 1 : (int)((nExtra + 31u) / 32u);
        for (int it = 0; it < iters; ++it) {
            // ---- the item of this thread: an entry of the list, or (last warp) an extra ------------------------------------
            uint32_t ent = 0;
            int slot = -1;
            bool isExtra = false;
            if (warp < RK_ENTRIES / 32) {
                if (haveEntry) { ent = myEnt; slot = tid; }
            } else {
                const uint32_t x = (uint32_t)it * 32u + (uint32_t)lane;
                if (x < nExtra) { ent = sh.extra[x]; slot = RK_ENTRIES + (int)x; isExtra = true; }
            }
            uint32_t statMask = 0, statMarked = 0, dynPairMask = 0;
            if (slot >= 0) {
                const int64_t s = ent >> 5;
                const int b = (int)(ent & 31u);
                const int64_t g = (int64_t)b * n + s;
                bool finished;
                unsigned long long keyPrev;
                float T;
                const StrikeEvent ev = strike_event(w, k, s, &finished, &keyPrev, &T);
                float tBest = __int_as_float(0x7fc00000);
                sh.ent[slot] = ent;
                sh.T[slot] = tBest;
                if (finished) {
                    // no event (billiard_search.cpp:1464-1467) or too many (:1376-1379): the strike ends in layer k - 1; its first
                    // dynamic ball's thread records it
                    if (!isExtra && b == __ffs(ev.dynNew) - 1) {
                        mark_finished(p, s, keyPrev == ~0ull ? F_NOEVENT : F_OVERFLOW, k - 1, T);
                        keyStale[s] = ~0ull;
                    }
                } else {
                    const BallRec me = new_state(old, n, s, b, ev, pc, sRails);
                    const bool leader = doApply && !isExtra && b == ev.ball;
                    if (leader) {
                        // ---- the strike's bookkeeping (System::append, type.cpp:467-531) -------------------------------------
                        w.maskDyn[cur][s] = ev.dynNew; w.maskRest[cur][s] = ev.restNew;
                        Tcur[s] = ev.newT;
                        keyStale[s] = ~0ull;
                        // A ball that came to rest or was potted in the PREVIOUS event is in layer k - 1's buffer only: bring its
                        // position and meta word over, so that balls that do not move are valid in both buffers (the other
                        // buffer could not be written while that layer was still being read)
                        const uint32_t stoppedBefore = w.stopped[prev][s];
                        if (stoppedBefore != 0xffu && !((ev.dynNew >> stoppedBefore) & 1u)) {
                            const int64_t gs = (int64_t)stoppedBefore * n + s;
                            neu.pos[gs] = old.pos[gs];
                            neu.meta[gs] = old.meta[gs];
                        }
                        const bool stays = ev.type == E_RAIL_COLLISION || ev.type == E_ROLLING || ev.type == E_BALL_COLLISION;
                        w.stopped[cur][s] = stays ? 0xffu : (uint32_t)ev.ball;
                        if (p.out.eventLog) {
                            int a = ev.ball, b2 = ev.partner;
                            if (ev.type == E_BALL_COLLISION && a > b2) { const int t = a; a = b2; b2 = t; }
                            if (ev.type == E_REST || ev.type == E_ROLLING) b2 = -1;
                            p.out.eventLog[s * 32 + (k - 1)] = make_int4(ev.type, a, b2, __float_as_int(ev.newT));
                        }
                        if (ev.dynNew == 0) mark_finished(p, s, F_COMPLETED, k, ev.newT);
                        else if (doItems) sh.appendLive[atomicAdd(&sh.nLiveAppend, 1u)] = (uint32_t)s;
                    }
                    const bool dynamicNow = ((ev.dynNew >> b) & 1u) != 0;
                    if (doApply) store_ball_to(neu, g, me);
                    if (dynamicNow && doItems) {
                        if (doApply) {
                            sh.appendDyn[atomicAdd(&sh.nDynAppend, 1u)] = ent;
                            offer_own(keyCur + s, me, b);   // nextBallInRest :1590, nextRolling :1575
                        }
                        sh.pos[slot] = make_float2(me.pos.x, me.pos.y); sh.vel[slot] = make_float2(me.vel.x, me.vel.y);
                        sh.vn[slot] = make_float2(me.vn.x, me.vn.y); sh.acc[slot] = make_float2(me.acc.x, me.acc.y);
                        // ---- stage 1 of items_kernel on the ball's layer-k state --------------------------------------------------
                        const v2 pos = me.pos, vel = me.vel, dn = me.vn, acc = me.acc;
                        const int lastRail = me.lastRail;
                        const uint32_t restMask = ev.restNew, dynMask = ev.dynNew;
                        unsigned long long keyNow = __ldcg(keyCur + s);   // rest / rolling offers of the strike so far (any value is a bound)
                        dynPairMask = dynMask & ~((2u << b) - 1u);
                        if (!is_zero(vel)) {
                            uint32_t railMask = 0, pocketMask = 0;
                            const bool reuse = prune && me.kind == K_MOVING && R <= 24;
                            if (reuse) {
                                const uint32_t c2 = w.lineCand[g];
                                railMask = c2 & 0xffffffu;
                                pocketMask = c2 >> 24;
                            } else {
                                const float speedSq = dot(vel, vel);
                                for (int r = 0; r < R; ++r) {
                                    const RailGeom rg = sRails[r];
                                    const float detDV = vel.x * rg.dy - vel.y * rg.dx;
                                    const float nT = (rg.dx * pos.y - rg.dy * pos.x) + sDetQV[r];
                                    bool cand = r != lastRail && detDV != 0.0f;
                                    if (prune) {
                                        const bool behind = nT * detDV < 0.0f && fabsf(nT) > 1e-10f * fabsf(detDV);
                                        const float lenSq = sRailLenSq[r];
                                        const float nU = (pos.x - rg.sx) * vel.y - (pos.y - rg.sy) * vel.x;   // u = nU / -detDV
                                        const float den = -detDV, sgnU = den < 0.0f ? -nU : nU, aden = fabsf(den);
                                        const bool wellCond = lenSq > 0.0f && detDV * detDV >= 4e-4f * (speedSq * lenSq);
                                        const bool outside = wellCond && (sgnU < -0.05f * aden || sgnU > 1.05f * aden);
                                        cand = cand && !behind && !outside;
                                    }
                                    if (cand) railMask |= 1u << r;
                                }
                                for (int kk = 0; kk < P; ++kk) {
                                    const PocketGeom pg = sPockets[kk];
                                    const v2 delta = pos - mk(pg.x, pg.y);
                                    const float qa = 1;
                                    const float qb = 2 * dot(dn, delta);
                                    const float qc = dot(delta, delta) - pg.r * pg.r;
                                    const float disc = qb * qb - 4 * qa * qc;
                                    const float slack = prune ? 1e-5f * (qb * qb + 4.0f * (dot(delta, delta) + pg.r * pg.r)) + 1.0f : 0.0f;
                                    if (!(disc < -slack) && !(prune && qb > 0.0f && qc > 0.0f)) pocketMask |= 1u << kk;
                                }
                                if (prune && R <= 24) w.lineCand[g] = railMask | (pocketMask << 24);
                            }
                            const float len1 = length(vel);
                            const float accLen = length(acc);
                            unsigned long long best = ~0ull;
                            while (railMask) {
                                const int r = __ffs(railMask) - 1;
                                railMask &= railMask - 1;
                                const RailGeom rg = sRails[r];
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
                                const int kk = __ffs(pocketMask) - 1;
                                pocketMask &= pocketMask - 1;
                                if (reuse) {
                                    const PocketGeom pg = sPockets[kk];
                                    const v2 delta = pos - mk(pg.x, pg.y);
                                    if (2 * dot(dn, delta) > 0.0f && dot(delta, delta) - pg.r * pg.r > 0.0f) continue;
                                }
                                float t;
                                if (time_to_pocket(sPockets[kk], pos, dn, vel, acc, &t) && t == t) {
                                    const unsigned long long k2 = ((unsigned long long)ordered_bits(t) << 32) | make_code(b, CAT_POCKET, kk);
                                    best = k2 < best ? k2 : best;
                                }
                            }
                            if (best != ~0ull) atomicMin(keyCur + s, best);
                            keyNow = best < keyNow ? best : keyNow;
                            tBest = from_ordered_bits((uint32_t)(keyNow >> 32));
                            sh.T[slot] = tBest;
                            const Reach rc = reach_of(tBest, diameter);
                            const float reach = rc.fixed + reach_ball(rc, len1, accLen);
                            const float reachSq = reach * reach;
                            const float passMax = 0.99f * (len1 * len1) / (2.0f * accLen) - diameter - 5.0f;
                            const float passMaxSq = passMax > 0.0f ? passMax * passMax : -1.0f;
                            const float corridor = diameter + 0.5f;
                            // resting partners: their positions are those of layer k - 1, except the event ball that has just come to rest
                            const float2* posOf = old.pos + s;
                            const int restEv = (ev.has && ev.type == E_REST) ? ev.ball : -1;
                            float2 restEvPos = make_float2(0.f, 0.f);
                            if (restEv >= 0) restEvPos = rest_position(old, n, s, restEv, ev);
                            uint32_t candMask = 0;
                            for (int o0 = 0; o0 < NB; o0 += PARTNER_BATCH) {
                                float2 qq[PARTNER_BATCH];
                                const float2* src = posOf + (int64_t)o0 * n;
                                #pragma unroll
                                for (int q = 0; q < PARTNER_BATCH; ++q) {
                                    qq[q] = (o0 + q < NB && ((restMask >> (o0 + q)) & 1u)) ? *src : make_float2(0.f, 0.f);
                                    if (o0 + q == restEv) qq[q] = restEvPos;
                                    src += n;
                                }
                                #pragma unroll
                                for (int q = 0; q < PARTNER_BATCH; ++q) {
                                    const float cr = dn.x * (qq[q].y - pos.y) - dn.y * (qq[q].x - pos.x);
                                    if (!prune || !(fabsf(cr) > corridor)) candMask |= 1u << (o0 + q);
                                }
                            }
                            candMask &= restMask;
                            while (candMask) {
                                const int o = __ffs(candMask) - 1;
                                candMask &= candMask - 1;
                                const float2 q2 = o == restEv ? restEvPos : posOf[(int64_t)o * n];
                                const v2 toPoint = mk(q2.x, q2.y) - pos;
                                const float factor = dot(dn, toPoint) / len1;
                                const v2 dv = toPoint - (vel * factor);
                                const float distSq = dot(toPoint, toPoint);
                                const float dist = factor < 0 ? distSq : dot(dv, dv);
                                if (!(dist > dd)) {
                                    bool skip = false;
                                    if (prune) {
#if PRUNE_REST_PASS_ONLY_V
                                        if (distSq <= passMaxSq) skip = distSq > reachSq;
#else
                                        skip = distSq > reachSq;
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
                }
            }
            push_mask(statMask & ~statMarked, (uint32_t)(slot & 0xff) << 6, qPairs, &cnt.pairs);
            if (prune == 2) push_mask(statMarked, ((uint32_t)(slot & 0xff) << 6) | 0x8000u, qPairs, &cnt.pairs);
            push_mask(dynPairMask, (uint32_t)(slot & 0xff) << 5, qDyn, &cnt.dynPairs);
        }
        __syncthreads();

        // ---- stage 1b: moving-vs-moving pre-checks, one queued pair per thread; the partner's layer-k state is derived here --
        {
            const int nd = cnt.dynPairs;
            for (int i0 = 0; i0 < nd; i0 += RK_BLOCK) {
                const int i = i0 + tid;
                bool pass = false;
                uint32_t code = 0;
                if (i < nd) {
                    const uint32_t c2 = qDyn[i];
                    const int slot = (int)(c2 >> 5);
                    const uint32_t en = sh.ent[slot];
                    const int o = (int)(c2 & 31u);
                    const int64_t s = en >> 5;
                    bool finished; unsigned long long keyPrev; float T;
                    const StrikeEvent ev = strike_event(w, k, s, &finished, &keyPrev, &T);
                    const BallRec other = new_state(old, n, s, o, ev, pc, sRails);
                    const float2 x1 = sh.pos[slot], y1 = sh.vel[slot], n1 = sh.vn[slot], z1 = sh.acc[slot];
                    const v2 p1 = mk(x1.x, x1.y), v1 = mk(y1.x, y1.y), p2 = other.pos, w2 = other.vel;
                    const v2 v1n = mk(n1.x, n1.y), w2n = other.vn;
                    bool outOfReach = false;
                    if (prune) {
                        const Reach rc = reach_of(sh.T[slot], diameter);
                        const v2 a1 = mk(z1.x, z1.y), a2 = other.acc;
                        const float reach = rc.fixed + reach_ball(rc, length(v1), length(a1)) + reach_ball(rc, length(w2), length(a2));
                        const v2 dA = (a1 - a2) * (1.0f / 1000.0f), dV = (v1 - w2) * (1.0f / 1000.0f), dP = p2 - p1;
                        const float aa = dot(dA, dA), vv = dot(dV, dV), cr = dA.x * dV.y - dA.y * dV.x;
                        outOfReach = 0.25f * aa >= PRUNE_MIN_LEADING && vv > 0.0f && cr * cr >= PRUNE_MIN_SIN_SQ * (aa * vv) &&
                                     !is_zero(v1) && !is_zero(w2) && dot(dP, dP) > reach * reach;
                        if (outOfReach) {
                            float co[5];
                            collision_coefficients(a1, v1, p1, a2, w2, p2, diameter, co);
                            const float t1 = co[1] * co[1] * co[1], t2 = 4.0f * co[0] * co[1] * co[2], t3 = 8.0f * co[0] * co[0] * co[3];
                            outOfReach = fabsf((t1 - t2) + t3) >= PRUNE_MIN_QREL * (fabsf(t1) + fabsf(t2) + fabsf(t3));
                        }
                    }
                    if (!outOfReach || prune == 2) {
                        #pragma unroll 1
                        for (int turn = 0; turn < 2 && !pass; ++turn) {
                            const v2 va = turn ? w2 : v1, pa = turn ? p2 : p1, van = turn ? w2n : v1n;
                            const v2 vb = turn ? v1 : w2, pb = turn ? p1 : p2, vbn = turn ? v1n : w2n;
                            if (is_zero(va)) continue;                                  // :287
                            if (dot(vb, vb) > 0) pass = collision_precheck_moving(va, pa, van, vb, pb, vbn, diameter);
                            else pass = !(point_to_half_line_sq(pa, va, pb) > diameter * diameter);   // :318-324
                        }
                    }
                    code = ((uint32_t)slot << 6) | 32u | (uint32_t)o;
                    if (pass && outOfReach) { ++skipped; code |= 0x8000u; }
                    else if (outOfReach && prune != 2) ++skipped;
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

        // ---- stage 2 + the lists of the next round: one reservation per tile --------------------------------------------------
        {
            const int cQ = cnt.pairs;
            if (tid == 0) {
                if (cQ > 0) sh.jobBase = atomicAdd(jobCount, (uint32_t)cQ);
                if (sh.nDynAppend) sh.dynBase = atomicAdd(&w.counts[4 * k + 1], sh.nDynAppend);
                if (sh.nLiveAppend) sh.liveBase = atomicAdd(&w.counts[4 * k + 0], sh.nLiveAppend);
            }
            __syncthreads();
            for (uint32_t i = tid; i < sh.nDynAppend; i += RK_BLOCK) dynNext[sh.dynBase + i] = sh.appendDyn[i];
            for (uint32_t i = tid; i < sh.nLiveAppend; i += RK_BLOCK) liveNext[sh.liveBase + i] = sh.appendLive[i];
            for (int iq = tid; iq < cQ; iq += RK_BLOCK) {
                const uint32_t c2 = qPairs[iq];
                const int slot = (int)((c2 >> 6) & 0xffu);
                const uint32_t en = sh.ent[slot];
                const int o = (int)(c2 & 31u);
                const int64_t s = en >> 5;
                const bool partnerDynamic = (c2 & 32u) != 0;
                bool finished; unsigned long long keyPrev; float T;
                const StrikeEvent ev = strike_event(w, k, s, &finished, &keyPrev, &T);
                v2 p2, v2o, a2o;
                if (partnerDynamic) {
                    const BallRec other = new_state(old, n, s, o, ev, pc, sRails);
                    p2 = other.pos; v2o = other.vel; a2o = other.acc;
                } else {
                    const float2 q = rest_position(old, n, s, o, ev);
                    p2 = mk(q.x, q.y); v2o = mk(0.f, 0.f); a2o = rest_acc(pc);
                }
                const float2 x1 = sh.pos[slot], y1 = sh.vel[slot], z1 = sh.acc[slot];
                float co[5];
                collision_coefficients(mk(z1.x, z1.y), mk(y1.x, y1.y), mk(x1.x, x1.y), a2o, v2o, p2, diameter, co);
                const uint64_t j = (uint64_t)sh.jobBase + iq;
                if (j < w.jobCap) {
                    #pragma unroll
                    for (int q = 0; q < 5; ++q) w.jobCo[(size_t)q * w.jobCap + j] = co[q];
                    w.jobCode[j] = (en << 6) | (c2 & 63u) | ((c2 & 0x8000u) ? JOB_MARK : 0u);
                } else {
                    w.counts[4 * ROUNDS + 3] = 1u;  // queue overflow: the host grows the queue and re-runs the launch
                }
            }
        }
        __syncthreads();
    }
    if (p.out.counters && skipped) atomicAdd(&p.out.counters[5], skipped);
}

__global__ void __launch_bounds__(RK_BLOCK, RK_MIN_BLOCKS) round_kernel(const LaunchParams* __restrict__ pp, PipeWorkspace w, int k) {
    extern __shared__ uint16_t smemQ[];
    __shared__ RoundShared sh;
    if (w.pdlEarly) cudaTriggerProgrammaticLaunchCompletion();
    {
        const int tid = threadIdx.x;
        for (int i = tid; i < w.nRails; i += RK_BLOCK) {
            const RailGeom rg = w.table->rails[i];
            sh.rails[i] = rg;
            sh.detQV[i] = rg.sx * rg.dy - rg.sy * rg.dx;
            const float used = fabsf(rg.dx != 0.0f ? rg.dx : rg.dy);
            sh.railLenSq[i] = used >= 5.0f ? rg.dx * rg.dx + rg.dy * rg.dy : -1.0f;
        }
        for (int i = tid; i < w.nPockets; i += RK_BLOCK) sh.pockets[i] = w.table->pockets[i];
    }
    cudaGridDependencySynchronize();
    __syncthreads();
    round_body(sh, smemQ, *pp, w, k, blockIdx.x, gridDim.x);
}
