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
    const int R = p.table->nRails, P = p.table->nPockets;
    for (int i = threadIdx.x; i < R; i += blockDim.x) sh.rails[i] = p.table->rails[i];
    for (int i = threadIdx.x; i < P; i += blockDim.x) sh.pockets[i] = p.table->pockets[i];
    __syncthreads();

    const PhysConst pc = c_phys;
    const float diameter = p.table->diameter;
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
                if (type == p.rules->cueType) {
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
            if (lane == 0) p.out.traceTimes[cand * 32] = T;
            if (lane < NB) {
                float* r = p.out.traceRec + ((size_t)cand * 32 * NB + lane) * 10;
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
                if (lane == 0) p.out.traceTimes[cand * 32 + nEvents] = T;
                if (lane < NB) {
                    float* r = p.out.traceRec + (((size_t)cand * 32 + nEvents) * NB + lane) * 10;
                    r[0] = (float)kind; r[1] = pos.x; r[2] = pos.y; r[3] = vel.x; r[4] = vel.y;
                    r[5] = rolling ? 1.f : 0.f; r[6] = velBefore.x; r[7] = velBefore.y; r[8] = (float)cause; r[9] = -1.f;
                    if (kind == K_ABSENT) for (int k = 0; k < 10; ++k) r[k] = -1.f;
                }
            }
        }

        // ---- epilogue: rules + simulationCost (billiard_search.cpp:1388-1405, 1473-1532) ------------------
        const unsigned pottedMask = __ballot_sync(FULL, kind == K_POTTING);
        const unsigned cuePotted = __ballot_sync(FULL, kind == K_POTTING && type == p.rules->cueType);
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
                if ((pottedMask >> b) & 1u) pottingScore += p.rules->score[tb];
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


void upload_constants(cudaStream_t stream) {
    PhysConst pc = phys_const();
    cudaMemcpyToSymbolAsync(c_phys, &pc, sizeof(PhysConst), 0, cudaMemcpyHostToDevice, stream);
    pipeline_upload_constants(stream);
    cudaStreamSynchronize(stream);
}

// The traced strikes (bcu_simulate_trace, bcu_trace_strikes): one warp per strike, full per-layer records.
cudaError_t launch_trace(const LaunchParams& p, cudaStream_t stream) {
    if (p.nCand <= 0) return cudaSuccess;
    const int64_t ctas = (p.nCand + WARPS_PER_CTA - 1) / WARPS_PER_CTA;
    simulate_candidates<true><<<(unsigned)(ctas < 148 * 8 ? ctas : 148 * 8), WARPS_PER_CTA * 32, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace bcu
