// K1 simulate_candidates: one warp simulates one candidate strike (north_star).
//
//   * lane l owns ball l (state in registers, mirrored in shared memory for pair access);
//   * every event step the warp enumerates  D x (rest, rolling, R rails, P pockets, N partners)
//     items across its lanes (D = dynamic balls), ball-pair items that survive the reference's
//     geometric pre-checks are stream-compacted (ballot + popc) into a per-warp job queue so the
//     FP64 quartic runs on densely packed lanes;
//   * the next event is the warp-min of (time, evaluation order) via __reduce_min_sync;
//   * the state is advanced in closed form by the owning lanes; bookkeeping follows
//     node::System::append (reference type.cpp:467-531);
//   * epilogue: rules (valid end state), simulationCost, potted mask, final positions, event log.
//
// Reference path: billiard_search.cpp:1358-1471 (simulate), :1550-1573 (nextEvent),
// :1957-1992 (createLayer), :1473-1532 (simulationCost); billiard_physics.cpp throughout.
// Compiled with -fmad=false: all arithmetic is un-contracted IEEE.
#include <cuda_runtime.h>

#include "bcu_kernels.cuh"
#include "bcu_physics.cuh"

namespace bcu {

__constant__ TableGeom c_table;
__constant__ RulesData c_rules;
__constant__ PhysConst c_phys;

constexpr int WARPS_PER_CTA = 8;
constexpr int QUEUE_CAP = 64;

struct WarpShared {
    float px[MAX_BALLS], py[MAX_BALLS];
    float vx[MAX_BALLS], vy[MAX_BALLS];
    float vnx[MAX_BALLS], vny[MAX_BALLS];
    float ax[MAX_BALLS], ay[MAX_BALLS];
    uint8_t rolling[MAX_BALLS];
    int8_t lastRail[MAX_BALLS];
    uint8_t dynList[MAX_BALLS];
    // quartic job queue
    float qco[QUEUE_CAP][5];
    uint32_t qcode[QUEUE_CAP];
};

struct CtaShared {
    RailGeom rails[MAX_RAILS];
    PocketGeom pockets[MAX_POCKETS];
    WarpShared w[WARPS_PER_CTA];
};

// order-preserving float -> uint32 (works for negative values too)
__device__ __forceinline__ uint32_t ordered_bits(float t) {
    uint32_t u = __float_as_uint(t + 0.0f);  // -0 -> +0
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float from_ordered_bits(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

enum Category : uint32_t { CAT_REST = 0, CAT_ROLLING = 1, CAT_STATIC_PAIR = 2, CAT_DYNAMIC_PAIR = 3, CAT_RAIL = 4, CAT_POCKET = 5 };
__device__ __forceinline__ uint32_t make_code(uint32_t ball, uint32_t cat, uint32_t partner) {
    return (ball << 16) | (cat << 8) | partner;
}

struct Best {
    uint32_t tbits;
    uint32_t code;
};
__device__ __forceinline__ void consider(Best& b, float t, uint32_t code) {
    if (t != t) return;  // a NaN time never wins the reference's strict '<' (billiard_search.cpp:2002-2013)
    uint32_t tb = ordered_bits(t);
    if (tb < b.tbits || (tb == b.tbits && code < b.code)) { b.tbits = tb; b.code = code; }
}

template <bool TRACE>
__global__ void __launch_bounds__(WARPS_PER_CTA * 32) simulate_candidates(LaunchParams p) {
    __shared__ CtaShared sh;
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int warpInCta = threadIdx.x >> 5;
    WarpShared& ws = sh.w[warpInCta];

    // stage the table geometry (same for every candidate)
    const int R = c_table.nRails, P = c_table.nPockets;
    for (int i = threadIdx.x; i < R; i += blockDim.x) sh.rails[i] = c_table.rails[i];
    for (int i = threadIdx.x; i < P; i += blockDim.x) sh.pockets[i] = c_table.pockets[i];
    __syncthreads();

    const PhysConst pc = c_phys;
    const float diameter = c_table.diameter;
    const int NB = p.states.nBalls;
    const int64_t warpsTotal = (int64_t)gridDim.x * WARPS_PER_CTA;
    unsigned long long cEvents = 0, cPairs = 0, cQuartics = 0, cRails = 0, cPockets = 0;

    for (int64_t cand = (int64_t)blockIdx.x * WARPS_PER_CTA + warpInCta; cand < p.nCand; cand += warpsTotal) {
        // ---- decode the candidate (K3 expand_level is this prologue) --------------------------------
        int sIdx;
        float cvx, cvy;
        uint64_t searchCost, baseCost;
        if (p.src.gridDirs > 0) {
            const int64_t gid = cand + p.src.firstCandidate;  // global candidate id of this shard's entry
            const int64_t perState = (int64_t)p.src.gridDirs * p.src.gridSpeeds;
            sIdx = (int)(gid / perState);
            const int rem = (int)(gid - (int64_t)sIdx * perState);
            const int k = rem / p.src.gridSpeeds, j = rem - k * p.src.gridSpeeds;
            const float sp = p.src.speedTable[j];
            cvx = sp * p.src.dirTable[2 * k];
            cvy = sp * p.src.dirTable[2 * k + 1];
            searchCost = p.src.gridSearchCost;
            baseCost = p.src.stateBaseCost ? p.src.stateBaseCost[sIdx] : 0ull;
        } else {
            sIdx = p.src.stateIdx[cand];
            cvx = p.src.vx[cand];
            cvy = p.src.vy[cand];
            searchCost = p.src.searchCost[cand];
            baseCost = p.src.baseCost[cand];
        }

        // ---- layer 0: toInputLayer (billiard_search.cpp:760-795) + System::append of it -----------------
        int type = -1;
        v2 pos = mk(0.f, 0.f), vel = mk(0.f, 0.f), vn, acc;
        bool rolling = false;
        int kind = K_ABSENT;
        float nextRolling = -1.0f;  // < 0: no pending rolling event
        int lastRail = -1;
        if (lane < NB) {
            type = p.states.type[(int64_t)sIdx * NB + lane];
            if (type >= 0) {
                const float2 q = reinterpret_cast<const float2*>(p.states.pos)[(int64_t)sIdx * NB + lane];
                pos = mk(q.x, q.y);
                if (type == c_rules.cueType) {
                    vel = mk(cvx, cvy);
                    kind = K_SHOT;
                    nextRolling = time_to_rolling(vel, pc);  // type.cpp:502-506
                } else {
                    kind = K_REST;
                }
            }
        }
        derive_ball(vel, rolling, pc, &vn, &acc);
        v2 velBefore = vel;
        int cause = -1;
        float T = 0.0f;
        int nEvents = 0;
        uint32_t flags = 0;

        if (TRACE) {
            if (lane == 0) p.out.traceTimes[0] = T;
            if (lane < NB) {
                float* r = p.out.traceRec + (size_t)lane * 10;
                r[0] = kind == K_ABSENT ? -1.f : (float)kind; r[1] = pos.x; r[2] = pos.y; r[3] = vel.x; r[4] = vel.y;
                r[5] = rolling ? 1.f : 0.f; r[6] = velBefore.x; r[7] = velBefore.y; r[8] = (float)cause; r[9] = -1.f;
                if (kind == K_ABSENT) for (int k = 1; k < 10; ++k) r[k] = -1.f;
            }
        }

        // ---- event loop (billiard_search.cpp:1370-1381) -----------------------------------------------
        for (;;) {
            const bool isDyn = kind == K_SHOT || kind == K_MOVING || kind == K_COLLISION || kind == K_RAIL;
            const unsigned dynMask = __ballot_sync(FULL, isDyn);
            const unsigned restMask = __ballot_sync(FULL, kind == K_REST);
            if (dynMask == 0) { flags |= F_COMPLETED; break; }
            const int D = __popc(dynMask);

            // publish state
            __syncwarp();
            ws.px[lane] = pos.x; ws.py[lane] = pos.y; ws.vx[lane] = vel.x; ws.vy[lane] = vel.y;
            ws.vnx[lane] = vn.x; ws.vny[lane] = vn.y; ws.ax[lane] = acc.x; ws.ay[lane] = acc.y;
            ws.rolling[lane] = rolling; ws.lastRail[lane] = (int8_t)lastRail;
            if (isDyn) ws.dynList[__popc(dynMask & ((1u << lane) - 1))] = (uint8_t)lane;
            __syncwarp();

            Best best;
            best.tbits = 0xffffffffu; best.code = 0xffffffffu;

            // own-ball items: rest (nextBallInRest) and rolling (nextRolling)
            if (isDyn) {
                consider(best, time_to_stop(acc, vel), make_code(lane, CAT_REST, 0));
                if (nextRolling >= 0.0f) consider(best, nextRolling, make_code(lane, CAT_ROLLING, 0));
            }
            // rails: D x R items
            for (int idx = lane; idx < D * R; idx += 32) {
                const int d = idx / R, r = idx - d * R;
                const int b = ws.dynList[d];
                if (ws.lastRail[b] == r) continue;
                float t;
                if (time_to_rail(mk(ws.px[b], ws.py[b]), mk(ws.vx[b], ws.vy[b]), mk(ws.ax[b], ws.ay[b]), sh.rails[r], &t))
                    consider(best, t, make_code(b, CAT_RAIL, r));
            }
            // pockets: D x P items
            for (int idx = lane; idx < D * P; idx += 32) {
                const int d = idx / P, k = idx - d * P;
                const int b = ws.dynList[d];
                float t;
                if (time_to_pocket(sh.pockets[k], mk(ws.px[b], ws.py[b]), mk(ws.vnx[b], ws.vny[b]), mk(ws.vx[b], ws.vy[b]),
                                   mk(ws.ax[b], ws.ay[b]), &t))
                    consider(best, t, make_code(b, CAT_POCKET, k));
            }
            // ball pairs: D x NB items -> pre-checks -> compacted quartic jobs
            int qCount = 0;  // warp-uniform
            const int nPairItems = D * NB;
            for (int base = 0; base < nPairItems || qCount > 0; base += 32) {
                const int idx = base + lane;
                bool job = false;
                float co[5];
                uint32_t code = 0;
                if (idx < nPairItems) {
                    const int d = idx / NB, o = idx - d * NB;
                    const int b = ws.dynList[d];
                    const bool oStatic = (restMask >> o) & 1u, oDyn = (dynMask >> o) & 1u;
                    if (o != b && (oStatic || (oDyn && b < o))) {
                        const v2 p1 = mk(ws.px[b], ws.py[b]), v1 = mk(ws.vx[b], ws.vy[b]);
                        const v2 p2 = mk(ws.px[o], ws.py[o]), w2 = mk(ws.vx[o], ws.vy[o]);
                        bool pass = collision_precheck(v1, p1, w2, p2, diameter);
                        if (!pass && oDyn) pass = collision_precheck(w2, p2, v1, p1, diameter);  // o's own turn
                        if (pass) {
                            collision_coefficients(mk(ws.ax[b], ws.ay[b]), v1, p1, mk(ws.ax[o], ws.ay[o]), w2, p2, diameter, co);
                            code = make_code(b, oStatic ? CAT_STATIC_PAIR : CAT_DYNAMIC_PAIR, o);
                            job = true;
                        }
                    }
                }
                const unsigned jm = __ballot_sync(FULL, job);
                if (job) {
                    const int slot = qCount + __popc(jm & ((1u << lane) - 1));
                    #pragma unroll
                    for (int k = 0; k < 5; ++k) ws.qco[slot][k] = co[k];
                    ws.qcode[slot] = code;
                }
                qCount += __popc(jm);
                __syncwarp();
                const bool lastRound = base + 32 >= nPairItems;
                if (qCount >= 32 || (lastRound && qCount > 0)) {
                    const int take = qCount < 32 ? qCount : 32;
                    const int first = qCount - take;  // pop from the tail
                    if (lane < take) {
                        float c5[5];
                        #pragma unroll
                        for (int k = 0; k < 5; ++k) c5[k] = ws.qco[first + lane][k];
                        float t;
                        if (collision_time_from_coefficients(c5, &t)) consider(best, t, ws.qcode[first + lane]);
                    }
                    if (p.out.counters) cQuartics += (lane == 0) ? take : 0;
                    qCount = first;
                    __syncwarp();
                }
            }
            if (p.out.counters && lane == 0) { cPairs += nPairItems; cRails += D * R; cPockets += D * P; }

            // ---- warp-min over (time, evaluation order) ----------------------------------------------
            const uint32_t minT = __reduce_min_sync(FULL, best.tbits);
            const uint32_t minCode = __reduce_min_sync(FULL, best.tbits == minT ? best.code : 0xffffffffu);
            if (minT == 0xffffffffu) { flags |= F_NOEVENT; break; }  // billiard_search.cpp:1464-1467
            if (nEvents + 1 > MAX_EVENTS) { flags |= F_OVERFLOW; break; }  // :1376-1379
            nEvents++;

            const float dt = from_ordered_bits(minT);
            const int evBall = (int)(minCode >> 16);
            const uint32_t cat = (minCode >> 8) & 0xffu;
            const int evPartner = (int)(minCode & 0xffu);
            const int evType = cat == CAT_REST ? E_REST : cat == CAT_ROLLING ? E_ROLLING : cat == CAT_RAIL ? E_RAIL_COLLISION
                             : cat == CAT_POCKET ? E_POTTING : E_BALL_COLLISION;

            // ---- createLayer (billiard_search.cpp:1957-1992) -------------------------------------------
            const float newT = T + dt;
            const float deltaTime = newT - T;  // type.cpp:475
            const bool affected = (lane == evBall) || (evType == E_BALL_COLLISION && lane == evPartner);
            if (isDyn && !affected) {
                const v2 nv = advance_velocity(acc, vel, dt);
                pos = advance_position(acc, vel, dt, pos);
                vel = nv; velBefore = nv; kind = K_MOVING; cause = -1;
                derive_ball(vel, rolling, pc, &vn, &acc);
            } else if (affected) {
                const v2 npos = advance_position(acc, vel, dt, pos);
                const v2 nvel = advance_velocity(acc, vel, dt);
                if (evType == E_POTTING) {            // :1716-1746
                    pos = npos; velBefore = nvel; vel = mk(0.f, 0.f); rolling = false; kind = K_POTTING; cause = evPartner;
                } else if (evType == E_RAIL_COLLISION) {  // :1748-1786
                    const RailGeom& rg = sh.rails[evPartner];
                    pos = npos; velBefore = nvel; vel = rail_collision(nvel, mk(rg.nx, rg.ny), pc);
                    rolling = true; kind = K_RAIL; cause = evPartner;
                } else if (evType == E_REST) {        // :1788-1808
                    pos = npos; vel = mk(0.f, 0.f); velBefore = vel; rolling = false; kind = K_REST; cause = -1;
                } else if (evType == E_ROLLING) {     // :1810-1840
                    pos = npos; vel = nvel; velBefore = nvel; rolling = true; kind = K_MOVING; cause = -1;
                } else {                              // :1842-1905 createBallCollisionNodes
                    const int other = lane == evBall ? evPartner : evBall;
                    const v2 op = mk(ws.px[other], ws.py[other]), ov = mk(ws.vx[other], ws.vy[other]);
                    const v2 oa = mk(ws.ax[other], ws.ay[other]);
                    const bool oroll = ws.rolling[other];
                    const v2 opos = advance_position(oa, ov, dt, op);
                    const v2 ovel = advance_velocity(oa, ov, dt);
                    v2 o1, o2;
                    // ball1 = evBall, ball2 = evPartner (the response is symmetric under exchanging them)
                    if (lane == evBall) elastic_collision(npos, nvel, opos, ovel, rolling, oroll, pc, &o1, &o2);
                    else elastic_collision(opos, ovel, npos, nvel, oroll, rolling, pc, &o2, &o1);
                    pos = npos; velBefore = nvel; vel = o1; kind = K_COLLISION; cause = other;
                }
                derive_ball(vel, rolling, pc, &vn, &acc);
            }
            // ---- System::append bookkeeping (type.cpp:478-527) ---------------------------------------
            if (kind != K_ABSENT) {
                if (nextRolling >= 0.0f) { const float nr = nextRolling - deltaTime; nextRolling = nr > 0.0f ? nr : 0.0f; }
                if (kind == K_COLLISION) {
                    lastRail = -1;
                    nextRolling = (!rolling && !is_zero(vel)) ? time_to_rolling(vel, pc) : -1.0f;
                } else if (kind == K_REST || kind == K_POTTING) {
                    nextRolling = -1.0f;
                } else if (kind == K_MOVING) {
                    if (rolling) nextRolling = -1.0f;
                } else if (kind == K_RAIL) {
                    lastRail = cause; nextRolling = -1.0f;
                }
            }
            T = newT;
            if (p.out.counters) cEvents += (lane == 0);

            if (p.out.eventLog && lane == 0) {
                int a = evBall, b = evPartner;
                if (evType == E_BALL_COLLISION && a > b) { int t = a; a = b; b = t; }
                if (evType == E_REST || evType == E_ROLLING) b = -1;
                p.out.eventLog[cand * 32 + (nEvents - 1)] = make_int4(evType, a, b, __float_as_int(T));
            }
            if (TRACE) {
                if (lane == 0) p.out.traceTimes[nEvents] = T;
                if (lane < NB) {
                    float* r = p.out.traceRec + ((size_t)nEvents * NB + lane) * 10;
                    r[0] = (float)kind; r[1] = pos.x; r[2] = pos.y; r[3] = vel.x; r[4] = vel.y;
                    r[5] = rolling ? 1.f : 0.f; r[6] = velBefore.x; r[7] = velBefore.y; r[8] = (float)cause; r[9] = -1.f;
                    if (kind == K_ABSENT) for (int k = 0; k < 10; ++k) r[k] = -1.f;
                }
            }
        }

        // ---- epilogue: rules + simulationCost (billiard_search.cpp:1388-1405, 1473-1532) ------------------
        const unsigned pottedMask = __ballot_sync(FULL, kind == K_POTTING);
        const unsigned cuePotted = __ballot_sync(FULL, kind == K_POTTING && type == c_rules.cueType);
        const uint32_t expected = p.states.expectedMask[sIdx];
        const unsigned unexpected = __ballot_sync(FULL, kind == K_POTTING && !((expected >> (type & 31)) & 1u));
        uint64_t cost = ~0ull;
        if (flags & F_COMPLETED) {
            // validEndState (billiard_snooker.cpp:1128-1153)
            if (pottedMask != 0 && cuePotted == 0 && unexpected == 0) flags |= F_VALID;
            // potting score summed in ball-index order
            double pottingScore = 0.0;
            for (int b = 0; b < NB; ++b) {
                const int tb = __shfl_sync(FULL, type, b);
                if ((pottedMask >> b) & 1u) pottingScore += c_rules.score[tb];
            }
            const v2 v0 = mk(cvx, cvy);
            const float velocitySquared = dot(v0, v0);
            const double velocityCost = (double)(velocitySquared / (5000.0f * 5000.0f));
            const double pottingCost = 1.0 - (pottingScore < 1.0 ? pottingScore : 1.0);
            const double wp = 0.5 * pottingCost, wv = 0.5 * velocityCost;
            const double total = (wp + wv) < 1.0 ? (wp + wv) : 1.0;
            const double cap = 1.0 * (double)searchCost;
            cost = baseCost + (uint64_t)(total * cap);
        }
        if (lane == 0) {
            p.out.cost[cand] = cost;
            p.out.flags[cand] = flags | ((uint32_t)nEvents << 8);
            p.out.potted[cand] = pottedMask;
            p.out.finalTime[cand] = T;
        }
        if (p.out.finalPos && lane < NB)
            reinterpret_cast<float2*>(p.out.finalPos)[cand * NB + lane] = make_float2(pos.x, pos.y);
    }

    if (p.out.counters && lane == 0) {
        atomicAdd(&p.out.counters[0], cEvents);
        atomicAdd(&p.out.counters[1], cPairs);
        atomicAdd(&p.out.counters[2], cQuartics);
        atomicAdd(&p.out.counters[3], cRails);
        atomicAdd(&p.out.counters[4], cPockets);
    }
}


// =====================================================================================================================
// K1 v2: the batch kernel.  Same arithmetic as the kernel above (which is kept for the single-strike trace), but laid
// out for lane density, which is what bounds this path (ncu on v1: 8.7 of 32 lanes active per instruction):
//
//   * every warp keeps S candidate strikes ("slots") in flight; their table states live in shared memory
//     (structure-of-arrays per slot: float4 {pos, vel}, float4 {velNormed, acc}, per-ball flags);
//   * each round advances every slot by one event.  The item spaces of all S slots
//         rails   : sum_s D_s x R        pockets : sum_s D_s x P        ball pairs : sum_s D_s x N
//     are enumerated 32 at a time; a cheap first stage filters them and the survivors are stream-compacted
//     (ballot + popc) into per-warp queues, so the expensive second stages -- rail / pocket quadratics with their
//     double-precision sqrt, and above all the FP64 complex quartic -- run on densely packed lanes;
//   * item results go to the slot's 64-bit key  (ordered time bits << 32 | evaluation-order code)  with a shared-memory
//     atomicMin: the minimum is the reference's "first evaluated wins" choice (billiard_search.cpp:2002-2013);
//   * a slot that reaches its final layer is costed, written out and refilled from a global atomic candidate
//     counter, so all slots stay busy and long strikes do not hold back short ones.
// =====================================================================================================================
constexpr int V2_WARPS = 8;
constexpr int V2_QCAP = 64;

struct SlotState {
    float4 pv[MAX_BALLS];        // pos.xy, vel.xy
    float4 na[MAX_BALLS];        // velNormed.xy, acc.xy   (state::BallState, type.cpp:176-184)
    float nextRolling[MAX_BALLS];  // < 0: none  (System::_nextRolling, type.cpp:478-527)
    uint8_t kind[MAX_BALLS];
    uint8_t rolling[MAX_BALLS];
    int8_t lastRail[MAX_BALLS];
    int8_t type[MAX_BALLS];
    unsigned long long best;
    unsigned long long searchCost, baseCost;
    long long cand;              // candidate index within this launch; < 0: slot empty
    float T, cvx, cvy;
    int nEvents, sIdx;
    unsigned dynMask, restMask;
};

template <int S>
struct WarpSharedV2 {
    SlotState slot[S];
    uint16_t gq[S * MAX_BALLS];  // global dynamic-ball list of the round: slot << 8 | ball
    // quartic jobs
    float qco[V2_QCAP][5];
    uint32_t qcode[V2_QCAP];     // slot << 24 | ball << 16 | cat << 8 | partner
    // rail / pocket second-stage jobs
    float rlam[V2_QCAP];
    uint32_t rcode[V2_QCAP];     // slot << 16 | ball << 8 | rail
    uint32_t pcode[V2_QCAP];     // slot << 16 | ball << 8 | pocket
};

template <int S>
struct CtaSharedV2 {
    RailGeom rails[MAX_RAILS];
    PocketGeom pockets[MAX_POCKETS];
    WarpSharedV2<S> w[V2_WARPS];
};

__device__ __forceinline__ unsigned long long make_key(float t, uint32_t code) {
    return ((unsigned long long)ordered_bits(t) << 32) | code;
}
__device__ __forceinline__ void offer(SlotState& sl, float t, uint32_t code) {
    if (t != t) return;  // NaN never wins the strict '<'
    atomicMin(&sl.best, make_key(t, code));
}

// toInputLayer (billiard_search.cpp:760-795) + System::append of layer 0, for one slot; lane = ball.
// Returns true when the candidate has at least one dynamic ball.
__device__ __forceinline__ bool v2_load_candidate(const LaunchParams& p, SlotState& sl, long long cand, int lane, int NB,
                                                  const PhysConst& pc) {
    int sIdx;
    float cvx, cvy;
    uint64_t searchCost, baseCost;
    if (p.src.gridDirs > 0) {
        const int64_t gid = cand + p.src.firstCandidate;
        const int64_t perState = (int64_t)p.src.gridDirs * p.src.gridSpeeds;
        sIdx = (int)(gid / perState);
        const int rem = (int)(gid - (int64_t)sIdx * perState);
        const int k = rem / p.src.gridSpeeds, j = rem - k * p.src.gridSpeeds;
        const float sp = p.src.speedTable[j];
        cvx = sp * p.src.dirTable[2 * k];
        cvy = sp * p.src.dirTable[2 * k + 1];
        searchCost = p.src.gridSearchCost;
        baseCost = p.src.stateBaseCost ? p.src.stateBaseCost[sIdx] : 0ull;
    } else {
        sIdx = p.src.stateIdx[cand];
        cvx = p.src.vx[cand];
        cvy = p.src.vy[cand];
        searchCost = p.src.searchCost[cand];
        baseCost = p.src.baseCost[cand];
    }
    int type = -1, kind = K_ABSENT;
    v2 pos = mk(0.f, 0.f), vel = mk(0.f, 0.f), vn, acc;
    float nextRolling = -1.0f;
    if (lane < NB) {
        type = p.states.type[(int64_t)sIdx * NB + lane];
        if (type >= 0) {
            const float2 q = reinterpret_cast<const float2*>(p.states.pos)[(int64_t)sIdx * NB + lane];
            pos = mk(q.x, q.y);
            if (type == c_rules.cueType) {
                vel = mk(cvx, cvy);
                kind = K_SHOT;
                nextRolling = time_to_rolling(vel, pc);
            } else {
                kind = K_REST;
            }
        }
    }
    derive_ball(vel, false, pc, &vn, &acc);
    sl.pv[lane] = make_float4(pos.x, pos.y, vel.x, vel.y);
    sl.na[lane] = make_float4(vn.x, vn.y, acc.x, acc.y);
    sl.nextRolling[lane] = nextRolling;
    sl.kind[lane] = (uint8_t)kind;
    sl.rolling[lane] = 0;
    sl.lastRail[lane] = -1;
    sl.type[lane] = (int8_t)type;
    if (lane == 0) {
        sl.cand = cand; sl.searchCost = searchCost; sl.baseCost = baseCost;
        sl.T = 0.0f; sl.cvx = cvx; sl.cvy = cvy; sl.nEvents = 0; sl.sIdx = sIdx;
    }
    return __any_sync(0xffffffffu, kind == K_SHOT);
}

// rules + simulationCost + result write-out for one finished slot (billiard_search.cpp:1388-1405, 1473-1532;
// billiard_snooker.cpp:1128-1153); lane = ball.
__device__ __forceinline__ void v2_finish(const LaunchParams& p, SlotState& sl, uint32_t flags, int lane, int NB) {
    const unsigned FULL = 0xffffffffu;
    const int kind = lane < NB ? sl.kind[lane] : K_ABSENT;
    const int type = lane < NB ? sl.type[lane] : -1;
    const long long cand = sl.cand;
    const unsigned pottedMask = __ballot_sync(FULL, kind == K_POTTING);
    const unsigned cuePotted = __ballot_sync(FULL, kind == K_POTTING && type == c_rules.cueType);
    const uint32_t expected = p.states.expectedMask[sl.sIdx];
    const unsigned unexpected = __ballot_sync(FULL, kind == K_POTTING && !((expected >> (type & 31)) & 1u));
    uint64_t cost = ~0ull;
    if (flags & F_COMPLETED) {
        if (pottedMask != 0 && cuePotted == 0 && unexpected == 0) flags |= F_VALID;
        double pottingScore = 0.0;
        for (unsigned m = pottedMask; m; m &= m - 1) {   // ascending ball index, as the oracle sums
            const int b = __ffs(m) - 1;
            pottingScore += c_rules.score[sl.type[b]];
        }
        const v2 v0 = mk(sl.cvx, sl.cvy);
        const float velocitySquared = dot(v0, v0);
        const double velocityCost = (double)(velocitySquared / (5000.0f * 5000.0f));
        const double pottingCost = 1.0 - (pottingScore < 1.0 ? pottingScore : 1.0);
        const double wp = 0.5 * pottingCost, wv = 0.5 * velocityCost;
        const double total = (wp + wv) < 1.0 ? (wp + wv) : 1.0;
        const double cap = 1.0 * (double)sl.searchCost;
        cost = sl.baseCost + (uint64_t)(total * cap);
    }
    if (lane == 0) {
        p.out.cost[cand] = cost;
        p.out.flags[cand] = flags | ((uint32_t)sl.nEvents << 8);
        p.out.potted[cand] = pottedMask;
        p.out.finalTime[cand] = sl.T;
    }
    if (p.out.finalPos && lane < NB) {
        const float4 q = sl.pv[lane];
        reinterpret_cast<float2*>(p.out.finalPos)[cand * NB + lane] = make_float2(q.x, q.y);
    }
}

template <int S>
__global__ void __launch_bounds__(V2_WARPS * 32) simulate_candidates_v2(LaunchParams p, unsigned long long* nextCand) {
    extern __shared__ __align__(16) unsigned char smemRaw[];
    CtaSharedV2<S>& sh = *reinterpret_cast<CtaSharedV2<S>*>(smemRaw);
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned ltMask = (1u << lane) - 1;
    WarpSharedV2<S>& ws = sh.w[threadIdx.x >> 5];

    const int R = c_table.nRails, P = c_table.nPockets;
    for (int i = threadIdx.x; i < R; i += blockDim.x) sh.rails[i] = c_table.rails[i];
    for (int i = threadIdx.x; i < P; i += blockDim.x) sh.pockets[i] = c_table.pockets[i];
    __syncthreads();

    const PhysConst pc = c_phys;
    const float diameter = c_table.diameter;
    const int NB = p.states.nBalls;
    unsigned long long cEvents = 0, cPairs = 0, cQuartics = 0, cRails = 0, cPockets = 0;

    // take the next candidate that has something to simulate; candidates without a dynamic ball are final at once
    auto refill = [&](SlotState& sl) {
        for (;;) {
            unsigned long long c = 0;
            if (lane == 0) c = atomicAdd(nextCand, 1ull);
            c = __shfl_sync(FULL, c, 0);
            __syncwarp();
            if ((long long)c >= p.nCand) { if (lane == 0) sl.cand = -1; __syncwarp(); return; }
            const bool live = v2_load_candidate(p, sl, (long long)c, lane, NB, pc);
            __syncwarp();
            if (live) return;
            v2_finish(p, sl, F_COMPLETED, lane, NB);
            __syncwarp();
        }
    };
    #pragma unroll 1
    for (int s = 0; s < S; ++s) refill(ws.slot[s]);

    for (;;) {
        // ---- A: per-slot masks, global dynamic-ball list, per-ball items (rest, rolling) --------------------------
        int Dtot = 0;
        bool anyActive = false;
        #pragma unroll 1
        for (int s = 0; s < S; ++s) {
            SlotState& sl = ws.slot[s];
            if (sl.cand < 0) continue;
            anyActive = true;
            const int kind = lane < NB ? sl.kind[lane] : K_ABSENT;
            const bool isDyn = kind == K_SHOT || kind == K_MOVING || kind == K_COLLISION || kind == K_RAIL;
            const unsigned dynMask = __ballot_sync(FULL, isDyn);
            const unsigned restMask = __ballot_sync(FULL, kind == K_REST);
            if (lane == 0) { sl.dynMask = dynMask; sl.restMask = restMask; sl.best = ~0ull; }
            if (isDyn) ws.gq[Dtot + __popc(dynMask & ltMask)] = (uint16_t)((s << 8) | lane);
            Dtot += __popc(dynMask);
            __syncwarp();
            if (isDyn) {
                const float4 q = sl.pv[lane], a = sl.na[lane];
                offer(sl, time_to_stop(mk(a.z, a.w), mk(q.z, q.w)), make_code(lane, CAT_REST, 0));
                const float nr = sl.nextRolling[lane];
                if (nr >= 0.0f) offer(sl, nr, make_code(lane, CAT_ROLLING, 0));
            }
        }
        if (!anyActive) break;
        __syncwarp();

        // ---- B1: rails.  Stage 1: half-line x shifted segment (billiard_physics.cpp:725-753); stage 2: the quadratic ----
        {
            const int total = Dtot * R;
            int rc = 0;
            for (int base = 0; base < total || rc > 0; base += 32) {
                const int idx = base + lane;
                bool surv = false;
                float lam = 0.f;
                uint32_t code = 0;
                if (idx < total) {
                    const int q = idx / R, r = idx - q * R;
                    const int sb = ws.gq[q], s = sb >> 8, b = sb & 255;
                    const SlotState& sl = ws.slot[s];
                    if (sl.lastRail[b] != r) {
                        const float4 st = sl.pv[b];
                        const v2 vel = mk(st.z, st.w);
                        const RailGeom rg = sh.rails[r];
                        float t, u;
                        if (!is_zero(vel) && line_intersects_line(mk(st.x, st.y), vel, mk(rg.sx, rg.sy), mk(rg.dx, rg.dy), &t, &u) &&
                            !(t < 0.0f) && !(u < 0.0f || u > 1.0f)) {
                            surv = true; lam = t; code = (uint32_t)((s << 16) | (b << 8) | r);
                        }
                    }
                }
                const unsigned m = __ballot_sync(FULL, surv);
                if (surv) { const int at = rc + __popc(m & ltMask); ws.rlam[at] = lam; ws.rcode[at] = code; }
                rc += __popc(m);
                __syncwarp();
                if (rc >= 32 || (base + 32 >= total && rc > 0)) {
                    const int take = rc < 32 ? rc : 32, first = rc - take;
                    if (lane < take) {
                        const uint32_t c2 = ws.rcode[first + lane];
                        const float lambda = ws.rlam[first + lane];
                        const int s = c2 >> 16, b = (c2 >> 8) & 255, r = c2 & 255;
                        SlotState& sl = ws.slot[s];
                        const float4 st = sl.pv[b], a4 = sl.na[b];
                        const v2 vel = mk(st.z, st.w);
                        // physics::timeToCollisionWithRail (:181-238) from the intersection onwards
                        const float distance = length(lambda * vel);
                        const float velocity = length(vel);
                        const float a = -length(mk(a4.z, a4.w));
                        float x[2];
                        if (solve_quadratic(0.5f * a, velocity, -distance, x) != 0) offer(sl, x[0], make_code(b, CAT_RAIL, r));
                    }
                    rc = first;
                    __syncwarp();
                }
            }
            if (p.out.counters && lane == 0) cRails += total;
        }

        // ---- B2: pockets.  Stage 1: does the travel line reach the pocket circle (discriminant of :584-625); stage 2: rest ----
        {
            const int total = Dtot * P;
            int rc = 0;
            for (int base = 0; base < total || rc > 0; base += 32) {
                const int idx = base + lane;
                bool surv = false;
                uint32_t code = 0;
                if (idx < total) {
                    const int q = idx / P, k = idx - q * P;
                    const int sb = ws.gq[q], s = sb >> 8, b = sb & 255;
                    const SlotState& sl = ws.slot[s];
                    const float4 st = sl.pv[b], a4 = sl.na[b];
                    if (!is_zero(mk(st.z, st.w))) {
                        const PocketGeom pk = sh.pockets[k];
                        const v2 delta = mk(st.x, st.y) - mk(pk.x, pk.y);
                        const float qa = 1;
                        const float qb = 2 * dot(mk(a4.x, a4.y), delta);
                        const float qc = dot(delta, delta) - pk.r * pk.r;
                        const float disc = qb * qb - 4 * qa * qc;
                        if (!(disc < 0)) { surv = true; code = (uint32_t)((s << 16) | (b << 8) | k); }
                    }
                }
                const unsigned m = __ballot_sync(FULL, surv);
                if (surv) ws.pcode[rc + __popc(m & ltMask)] = code;
                rc += __popc(m);
                __syncwarp();
                if (rc >= 32 || (base + 32 >= total && rc > 0)) {
                    const int take = rc < 32 ? rc : 32, first = rc - take;
                    if (lane < take) {
                        const uint32_t c2 = ws.pcode[first + lane];
                        const int s = c2 >> 16, b = (c2 >> 8) & 255, k = c2 & 255;
                        SlotState& sl = ws.slot[s];
                        const float4 st = sl.pv[b], a4 = sl.na[b];
                        float t;
                        if (time_to_pocket(sh.pockets[k], mk(st.x, st.y), mk(a4.x, a4.y), mk(st.z, st.w), mk(a4.z, a4.w), &t))
                            offer(sl, t, make_code(b, CAT_POCKET, k));
                    }
                    rc = first;
                    __syncwarp();
                }
            }
            if (p.out.counters && lane == 0) cPockets += total;
        }

        // ---- B3: ball pairs.  Stage 1: the reference's geometric pre-checks + quantised coefficients; stage 2: quartic ----
        {
            const int total = Dtot * NB;
            int qCount = 0;
            for (int base = 0; base < total || qCount > 0; base += 32) {
                const int idx = base + lane;
                bool job = false;
                float co[5];
                uint32_t code = 0;
                if (idx < total) {
                    const int q = idx / NB, o = idx - q * NB;
                    const int sb = ws.gq[q], s = sb >> 8, b = sb & 255;
                    const SlotState& sl = ws.slot[s];
                    const bool oStatic = (sl.restMask >> o) & 1u, oDyn = (sl.dynMask >> o) & 1u;
                    if (o != b && (oStatic || (oDyn && b < o))) {
                        const float4 s1 = sl.pv[b], s2 = sl.pv[o];
                        const v2 p1 = mk(s1.x, s1.y), v1 = mk(s1.z, s1.w), p2 = mk(s2.x, s2.y), w2 = mk(s2.z, s2.w);
                        // the reference tests b -> o and, for a dynamic partner, o -> b on o's own turn; the quantised
                        // coefficients of both orders are identical, so one job covers both
                        bool pass = false;
                        #pragma unroll 1
                        for (int turn = 0; turn < (oDyn ? 2 : 1) && !pass; ++turn)
                            pass = turn == 0 ? collision_precheck(v1, p1, w2, p2, diameter) : collision_precheck(w2, p2, v1, p1, diameter);
                        if (pass) {
                            const float4 a1 = sl.na[b], a2 = sl.na[o];
                            collision_coefficients(mk(a1.z, a1.w), v1, p1, mk(a2.z, a2.w), w2, p2, diameter, co);
                            code = ((uint32_t)s << 24) | make_code(b, oStatic ? CAT_STATIC_PAIR : CAT_DYNAMIC_PAIR, o);
                            job = true;
                        }
                    }
                }
                const unsigned jm = __ballot_sync(FULL, job);
                if (job) {
                    const int at = qCount + __popc(jm & ltMask);
                    #pragma unroll
                    for (int k = 0; k < 5; ++k) ws.qco[at][k] = co[k];
                    ws.qcode[at] = code;
                }
                qCount += __popc(jm);
                __syncwarp();
                if (qCount >= 32 || (base + 32 >= total && qCount > 0)) {
                    const int take = qCount < 32 ? qCount : 32, first = qCount - take;
                    if (lane < take) {
                        float c5[5];
                        #pragma unroll
                        for (int k = 0; k < 5; ++k) c5[k] = ws.qco[first + lane][k];
                        const uint32_t c2 = ws.qcode[first + lane];
                        float t;
                        if (collision_time_from_coefficients(c5, &t)) offer(ws.slot[c2 >> 24], t, c2 & 0x00ffffffu);
                    }
                    if (p.out.counters) cQuartics += (lane == 0) ? take : 0;
                    qCount = first;
                    __syncwarp();
                }
            }
            if (p.out.counters && lane == 0) cPairs += total;
        }
        __syncwarp();

        // ---- C: apply the winning event of every slot (createLayer :1957-1992, System::append type.cpp:467-531) ------
        #pragma unroll 1
        for (int s = 0; s < S; ++s) {
            SlotState& sl = ws.slot[s];
            if (sl.cand < 0) continue;
            const unsigned long long key = sl.best;
            const int nEvents = sl.nEvents;
            uint32_t endFlags = 0;
            if (key == ~0ull) endFlags = F_NOEVENT;                       // billiard_search.cpp:1464-1467
            else if (nEvents + 1 > MAX_EVENTS) endFlags = F_OVERFLOW;     // :1376-1379
            if (endFlags == 0) {
                const float dt = from_ordered_bits((uint32_t)(key >> 32));
                const uint32_t minCode = (uint32_t)key;
                const int evBall = (int)(minCode >> 16);
                const uint32_t cat = (minCode >> 8) & 0xffu;
                const int evPartner = (int)(minCode & 0xffu);
                const int evType = cat == CAT_REST ? E_REST : cat == CAT_ROLLING ? E_ROLLING : cat == CAT_RAIL ? E_RAIL_COLLISION
                                 : cat == CAT_POCKET ? E_POTTING : E_BALL_COLLISION;
                const float T = sl.T;
                const float newT = T + dt;
                const float deltaTime = newT - T;  // type.cpp:475
                int kind = K_ABSENT;
                v2 pos = mk(0.f, 0.f), vel = mk(0.f, 0.f), vn = mk(0.f, 0.f), acc = mk(0.f, 0.f);
                bool rolling = false;
                float nextRolling = -1.0f;
                int lastRail = -1;
                if (lane < NB) {
                    kind = sl.kind[lane];
                    const float4 st = sl.pv[lane], a4 = sl.na[lane];
                    pos = mk(st.x, st.y); vel = mk(st.z, st.w); vn = mk(a4.x, a4.y); acc = mk(a4.z, a4.w);
                    rolling = sl.rolling[lane] != 0;
                    nextRolling = sl.nextRolling[lane];
                    lastRail = sl.lastRail[lane];
                    const bool isDyn = kind == K_SHOT || kind == K_MOVING || kind == K_COLLISION || kind == K_RAIL;
                    const bool affected = (lane == evBall) || (evType == E_BALL_COLLISION && lane == evPartner);
                    int cause = -1;
                    bool changed = false;
                    if (isDyn && !affected) {
                        const v2 nv = advance_velocity(acc, vel, dt);
                        pos = advance_position(acc, vel, dt, pos);
                        vel = nv; kind = K_MOVING; changed = true;
                    } else if (affected) {
                        const v2 npos = advance_position(acc, vel, dt, pos);
                        const v2 nvel = advance_velocity(acc, vel, dt);
                        if (evType == E_POTTING) {
                            pos = npos; vel = mk(0.f, 0.f); rolling = false; kind = K_POTTING; cause = evPartner;
                        } else if (evType == E_RAIL_COLLISION) {
                            const RailGeom& rg = sh.rails[evPartner];
                            pos = npos; vel = rail_collision(nvel, mk(rg.nx, rg.ny), pc);
                            rolling = true; kind = K_RAIL; cause = evPartner;
                        } else if (evType == E_REST) {
                            pos = npos; vel = mk(0.f, 0.f); rolling = false; kind = K_REST;
                        } else if (evType == E_ROLLING) {
                            pos = npos; vel = nvel; rolling = true; kind = K_MOVING;
                        } else {
                            const int other = lane == evBall ? evPartner : evBall;
                            const float4 os = sl.pv[other], oa4 = sl.na[other];
                            const bool oroll = sl.rolling[other] != 0;
                            const v2 opos = advance_position(mk(oa4.z, oa4.w), mk(os.z, os.w), dt, mk(os.x, os.y));
                            const v2 ovel = advance_velocity(mk(oa4.z, oa4.w), mk(os.z, os.w), dt);
                            // ball1 = evBall, ball2 = evPartner: one call site, operands swapped by select
                            const bool first = lane == evBall;
                            v2 oa, ob;
                            elastic_collision(first ? npos : opos, first ? nvel : ovel, first ? opos : npos, first ? ovel : nvel,
                                              first ? rolling : oroll, first ? oroll : rolling, pc, &oa, &ob);
                            pos = npos; vel = first ? oa : ob; kind = K_COLLISION; cause = other;
                        }
                        changed = true;
                    }
                    if (changed) derive_ball(vel, rolling, pc, &vn, &acc);
                    if (kind != K_ABSENT) {
                        if (nextRolling >= 0.0f) { const float nr = nextRolling - deltaTime; nextRolling = nr > 0.0f ? nr : 0.0f; }
                        if (kind == K_COLLISION) {
                            lastRail = -1;
                            nextRolling = (!rolling && !is_zero(vel)) ? time_to_rolling(vel, pc) : -1.0f;
                        } else if (kind == K_REST || kind == K_POTTING) {
                            nextRolling = -1.0f;
                        } else if (kind == K_MOVING) {
                            if (rolling) nextRolling = -1.0f;
                        } else if (kind == K_RAIL) {
                            lastRail = cause; nextRolling = -1.0f;
                        }
                    }
                }
                __syncwarp();  // partner states were read above; now publish
                if (lane < NB) {
                    sl.pv[lane] = make_float4(pos.x, pos.y, vel.x, vel.y);
                    sl.na[lane] = make_float4(vn.x, vn.y, acc.x, acc.y);
                    sl.nextRolling[lane] = nextRolling;
                    sl.kind[lane] = (uint8_t)kind;
                    sl.rolling[lane] = rolling;
                    sl.lastRail[lane] = (int8_t)lastRail;
                }
                if (lane == 0) {
                    sl.T = newT; sl.nEvents = nEvents + 1;
                    if (p.out.eventLog) {
                        int a = evBall, b = evPartner;
                        if (evType == E_BALL_COLLISION && a > b) { const int t = a; a = b; b = t; }
                        if (evType == E_REST || evType == E_ROLLING) b = -1;
                        p.out.eventLog[sl.cand * 32 + nEvents] = make_int4(evType, a, b, __float_as_int(newT));
                    }
                }
                if (p.out.counters) cEvents += (lane == 0);
                const bool stillDyn = kind == K_SHOT || kind == K_MOVING || kind == K_COLLISION || kind == K_RAIL;
                if (__ballot_sync(FULL, stillDyn) == 0) endFlags = F_COMPLETED;
            }
            __syncwarp();
            if (endFlags != 0) {
                v2_finish(p, sl, endFlags, lane, NB);
                __syncwarp();
                refill(sl);
            }
        }
        __syncwarp();
    }

    if (p.out.counters && lane == 0) {
        atomicAdd(&p.out.counters[0], cEvents);
        atomicAdd(&p.out.counters[1], cPairs);
        atomicAdd(&p.out.counters[2], cQuartics);
        atomicAdd(&p.out.counters[3], cRails);
        atomicAdd(&p.out.counters[4], cPockets);
    }
}

constexpr int V2_SLOTS = 4;

void upload_constants(const TableGeom& table, const RulesData& rules, cudaStream_t stream) {
    PhysConst pc = phys_const();
    cudaMemcpyToSymbolAsync(c_table, &table, sizeof(TableGeom), 0, cudaMemcpyHostToDevice, stream);
    cudaMemcpyToSymbolAsync(c_rules, &rules, sizeof(RulesData), 0, cudaMemcpyHostToDevice, stream);
    cudaMemcpyToSymbolAsync(c_phys, &pc, sizeof(PhysConst), 0, cudaMemcpyHostToDevice, stream);
    pipeline_upload_constants(table, rules, stream);
    cudaStreamSynchronize(stream);
}

int simulate_max_ctas(int device) {
    int sms = 0, perSm = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const size_t smem = sizeof(CtaSharedV2<V2_SLOTS>);
    cudaFuncSetAttribute(simulate_candidates_v2<V2_SLOTS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, simulate_candidates_v2<V2_SLOTS>, V2_WARPS * 32, smem);
    if (perSm < 1) perSm = 1;
    return sms * perSm;
}

cudaError_t launch_simulate(const LaunchParams& p, bool trace, int ctas, unsigned long long* nextCand, cudaStream_t stream) {
    if (p.nCand <= 0) return cudaSuccess;
    if (trace) {
        simulate_candidates<true><<<1, WARPS_PER_CTA * 32, 0, stream>>>(p);
        return cudaGetLastError();
    }
    // persistent grid: one resident wave; every warp pulls candidates from *nextCand until none are left
    const int64_t need = (p.nCand + V2_WARPS * V2_SLOTS - 1) / (V2_WARPS * V2_SLOTS);
    const int grid = (int)(need < ctas ? need : ctas);
    cudaError_t e = cudaMemsetAsync(nextCand, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    simulate_candidates_v2<V2_SLOTS><<<grid, V2_WARPS * 32, sizeof(CtaSharedV2<V2_SLOTS>), stream>>>(p, nextCand);
    return cudaGetLastError();
}

}  // namespace bcu
