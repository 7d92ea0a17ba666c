// plan_kernel: the reference's geometric planner (bcu_planner.cuh) over every state of the beam at once.
//
// One thread per (state, search pocket, first ball, first link, second ball, cushions of the second link) -- the first link being the direct one or one rail
// combination of one to three cushions (1 + 3 x rails-in-search variants) -- builds that first search node and enumerates
// everything below its child for the second ball depth first, emitting the strikes of every path that reaches the cue ball.
// The second ball is the fastest-running index, so the lanes of a warp share the first node (one converged instruction
// stream up to it) and fan out over its children; a first node that is not feasible retires its 16-32 lanes at once.  (With
// the first link as the unit, profiles/r02_plan_count_kernel_ncu_full.txt showed ONE active lane per warp: a feasible first
// node carries a whole subtree, its neighbours nothing.)  Integer / FP32 / a little FP64 scalar work with long
// divergent paths: not a throughput kernel; what matters is that the candidate lists of a whole beam never leave the device
// (the reference plans one state at a time on the host: ~1 700 expansions, 47 ms per state).
//
// The plan has to be deterministic (candidate ids break ties in the top-k), so it is made in two passes over the same
// enumeration: count per thread -> exclusive scan (per block, then over the block sums) -> write at the thread's offset.
#include <cuda_runtime.h>

#include "bcu_kernels.cuh"
#include "bcu_planner.cuh"

namespace bcu {

namespace {

constexpr int PLAN_BLOCK = 128;

struct CountEmit {
    uint32_t n = 0;
    __device__ void operator()(float, float, uint64_t, uint64_t) { ++n; }
};
struct WriteEmit {
    int64_t at;
    int32_t state;
    int32_t* stateIdx; float* vx; float* vy; uint64_t* searchCost; uint64_t* cost;
    __device__ void operator()(float x, float y, uint64_t sc, uint64_t c) {
        stateIdx[at] = state; vx[at] = x; vy[at] = y; searchCost[at] = sc; cost[at] = c;
        ++at;
    }
};

struct PlanWork {   // decode of a work item index
    int state, pocket;
    PlanFilter filter;
};
__device__ __forceinline__ bool plan_decode(const PlanTable& tb, int nStates, int nBalls, int64_t w, PlanWork* out) {
    const int variants = 1 + PLAN_BANK_INDIRECTION * tb.nSetOrder;
    const int64_t total = (int64_t)nStates * tb.nPockets * nBalls * variants * nBalls * (PLAN_BANK_INDIRECTION + 1);
    if (w >= total) return false;
    const int secondDepth = (int)(w % (PLAN_BANK_INDIRECTION + 1)); w /= (PLAN_BANK_INDIRECTION + 1);
    const int second = (int)(w % nBalls); w /= nBalls;
    const int v = (int)(w % variants); w /= variants;
    const int ball = (int)(w % nBalls); w /= nBalls;
    out->pocket = (int)(w % tb.nPockets);
    out->state = (int)(w / tb.nPockets);
    const int depth = v == 0 ? 0 : 1 + (v - 1) / tb.nSetOrder;
    const int combo = v == 0 ? 0 : (v - 1) % tb.nSetOrder;
    out->filter.ball = ball; out->filter.depth = depth; out->filter.combo = combo;
    out->filter.nextBall = second; out->filter.nextDepth = secondDepth;
    return true;
}

template <class Emit>
__device__ __forceinline__ void plan_item(const PlanTable& tb, const StateTable& states, const uint64_t* stateBase, const PlanWork& wk, Emit& emit) {
    PlanState st;
    st.nBalls = states.nBalls;
    st.pos = states.pos + (size_t)wk.state * states.nBalls * 2;
    st.type = states.type + (size_t)wk.state * states.nBalls;
    st.searchMask = states.expectedMask[wk.state];
    if ((st.searchMask >> tb.cueType) & 1u) return;   // isValidSearch (billiard_search.cpp:100-120): the cue ball is never searched for
    plan_root(tb, st, wk.pocket, stateBase ? stateBase[wk.state] : 0ull, phys_const(), emit, wk.filter);
}

// pass 0: strikes per work item and per block
__global__ void __launch_bounds__(PLAN_BLOCK) plan_count_kernel(const PlanTable* __restrict__ tbp, StateTable states, const uint64_t* stateBase,
                                                                 uint32_t* __restrict__ itemCount, uint32_t* __restrict__ blockSum) {
    __shared__ uint32_t warpSum[PLAN_BLOCK / 32];
    const PlanTable& tb = *tbp;
    const int64_t w = (int64_t)blockIdx.x * PLAN_BLOCK + threadIdx.x;
    PlanWork wk;
    CountEmit emit;
    if (plan_decode(tb, states.nStates, states.nBalls, w, &wk)) {
        plan_item(tb, states, stateBase, wk, emit);
        itemCount[w] = emit.n;
    }
    uint32_t v = emit.n;
    #pragma unroll
    for (int d = 16; d >= 1; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) == 0) warpSum[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t s = 0;
        for (int i = 0; i < PLAN_BLOCK / 32; ++i) s += warpSum[i];
        blockSum[blockIdx.x] = s;
    }
}

// exclusive scan of the block sums (one block; the plan of a beam is a few thousand blocks)
__global__ void __launch_bounds__(1024) plan_scan_kernel(const uint32_t* __restrict__ blockSum, int64_t nBlocks, unsigned long long* __restrict__ blockOffset,
                                                          unsigned long long* __restrict__ total) {
    __shared__ unsigned long long part[1024];
    const int64_t per = (nBlocks + 1023) / 1024;
    const int64_t lo = (int64_t)threadIdx.x * per, hi = lo + per < nBlocks ? lo + per : nBlocks;
    unsigned long long s = 0;
    for (int64_t i = lo; i < hi; ++i) s += blockSum[i];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long run = 0;
        for (int i = 0; i < 1024; ++i) { const unsigned long long t = part[i]; part[i] = run; run += t; }
        *total = run;
    }
    __syncthreads();
    unsigned long long run = part[threadIdx.x];
    for (int64_t i = lo; i < hi; ++i) { blockOffset[i] = run; run += blockSum[i]; }
}

// pass 1: the same enumeration, written at the item's offset
__global__ void __launch_bounds__(PLAN_BLOCK) plan_write_kernel(const PlanTable* __restrict__ tbp, StateTable states, const uint64_t* stateBase,
                                                                 const uint32_t* __restrict__ itemCount, const unsigned long long* __restrict__ blockOffset,
                                                                 int32_t* stateIdx, float* vx, float* vy, uint64_t* searchCost, uint64_t* cost) {
    __shared__ uint32_t sCount[PLAN_BLOCK];
    const PlanTable& tb = *tbp;
    const int64_t w = (int64_t)blockIdx.x * PLAN_BLOCK + threadIdx.x;
    PlanWork wk;
    const bool have = plan_decode(tb, states.nStates, states.nBalls, w, &wk);
    sCount[threadIdx.x] = have ? itemCount[w] : 0u;
    __syncthreads();
    if (!have || sCount[threadIdx.x] == 0) return;
    unsigned long long at = blockOffset[blockIdx.x];
    for (int i = 0; i < (int)threadIdx.x; ++i) at += sCount[i];
    WriteEmit emit{(int64_t)at, wk.state, stateIdx, vx, vy, searchCost, cost};
    plan_item(tb, states, stateBase, wk, emit);
}

}  // namespace

int64_t plan_work_items(const PlanTable& tb, int nStates, int nBalls) {
    return (int64_t)nStates * tb.nPockets * nBalls * (1 + PLAN_BANK_INDIRECTION * tb.nSetOrder) * nBalls * (PLAN_BANK_INDIRECTION + 1);
}

cudaError_t launch_plan_count(const PlanTable* dTable, const PlanTable& hTable, const StateTable& states, const uint64_t* stateBase,
                              uint32_t* itemCount, uint32_t* blockSum, unsigned long long* blockOffset, unsigned long long* total,
                              cudaStream_t stream) {
    const int64_t items = plan_work_items(hTable, states.nStates, states.nBalls);
    const int64_t blocks = (items + PLAN_BLOCK - 1) / PLAN_BLOCK;
    if (blocks == 0) return cudaMemsetAsync(total, 0, sizeof(unsigned long long), stream);
    plan_count_kernel<<<(unsigned)blocks, PLAN_BLOCK, 0, stream>>>(dTable, states, stateBase, itemCount, blockSum);
    plan_scan_kernel<<<1, 1024, 0, stream>>>(blockSum, blocks, blockOffset, total);
    return cudaGetLastError();
}

cudaError_t launch_plan_write(const PlanTable* dTable, const PlanTable& hTable, const StateTable& states, const uint64_t* stateBase,
                              const uint32_t* itemCount, const unsigned long long* blockOffset, int32_t* stateIdx, float* vx, float* vy,
                              uint64_t* searchCost, uint64_t* cost, cudaStream_t stream) {
    const int64_t items = plan_work_items(hTable, states.nStates, states.nBalls);
    const int64_t blocks = (items + PLAN_BLOCK - 1) / PLAN_BLOCK;
    if (blocks == 0) return cudaSuccess;
    plan_write_kernel<<<(unsigned)blocks, PLAN_BLOCK, 0, stream>>>(dTable, states, stateBase, itemCount, blockOffset, stateIdx, vx, vy, searchCost, cost);
    return cudaGetLastError();
}

int64_t plan_blocks(const PlanTable& tb, int nStates, int nBalls) { return (plan_work_items(tb, nStates, nBalls) + PLAN_BLOCK - 1) / PLAN_BLOCK; }

}  // namespace bcu
