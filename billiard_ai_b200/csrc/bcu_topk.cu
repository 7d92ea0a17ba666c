// K2 topk: block-level top-k on (cost, candidate id) among VALID candidates.
//
// The reference returns solutions in ascending cost (process::Comparator, process.hpp:19-25, and
// SearchNode::_cost, type.hpp:572); ties are broken here by the global candidate id so that the
// selection is deterministic and identical on every rank.
//
// Pass 1 (topk_select): each CTA compacts the valid keys of its slice into a 2048-entry bitonic sorter in shared memory,
// keeping a running best-1024.  Pass 2 (topk_rank_merge): the sorted per-CTA lists are merged by rank.
#include <cuda_runtime.h>

#include "bcu_kernels.cuh"

namespace bcu {

constexpr int TK_THREADS = 1024;
constexpr int TK_CAP = 1024;  // max k

struct Key {
    uint64_t cost, id;
};
__device__ __forceinline__ bool key_less(uint64_t c1, uint64_t i1, uint64_t c2, uint64_t i2) {
    return c1 < c2 || (c1 == c2 && i1 < i2);
}

// ascending bitonic sort of 2*TK_CAP keys held in shared memory
__device__ void bitonic_sort_2048(uint64_t* sc, uint64_t* si) {
    const int tid = threadIdx.x;
    for (int size = 2; size <= 2 * TK_CAP; size <<= 1) {
        for (int stride = size >> 1; stride > 0; stride >>= 1) {
            __syncthreads();
            const int pos = 2 * tid - (tid & (stride - 1));
            const bool up = ((pos & size) == 0);
            const uint64_t c1 = sc[pos], i1 = si[pos], c2 = sc[pos + stride], i2 = si[pos + stride];
            const bool lt = key_less(c2, i2, c1, i1);
            if (lt == up) { sc[pos] = c2; si[pos] = i2; sc[pos + stride] = c1; si[pos + stride] = i1; }
        }
    }
    __syncthreads();
}

// Pass 1.  Each CTA streams its slice of the candidates: VALID keys are compacted (ballot + prefix) into the upper half
// of the 2048-entry sorter, and only when that half could overflow (or at the end) the sorter runs, leaving the running
// best-1024 in the lower half.  With the usual few-percent share of valid strikes a CTA sorts once.
__global__ void __launch_bounds__(TK_THREADS) topk_select(const uint64_t* inCost, const uint32_t* inFlags, int64_t n,
                                                          int64_t firstCandidate, int k, uint64_t* outCost, uint64_t* outId,
                                                          int32_t* outValid) {
    __shared__ uint64_t sc[2 * TK_CAP];
    __shared__ uint64_t si[2 * TK_CAP];
    __shared__ int sWarp[TK_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    sc[tid] = ~0ull; si[tid] = ~0ull;
    sc[TK_CAP + tid] = ~0ull; si[TK_CAP + tid] = ~0ull;
    const int64_t per = (n + gridDim.x - 1) / gridDim.x;
    const int64_t lo = per * blockIdx.x;
    const int64_t hi = lo + per < n ? lo + per : n;
    int staged = 0, totalValid = 0;   // block-uniform
    for (int64_t base = lo; base < hi; base += TK_CAP) {
        const int64_t i = base + tid;
        const bool valid = i < hi && (inFlags[i] & F_VALID);
        const unsigned m = __ballot_sync(0xffffffffu, valid);
        if (lane == 0) sWarp[warp] = __popc(m);
        __syncthreads();
        int before = 0, cnt = 0;
        #pragma unroll
        for (int w = 0; w < TK_THREADS / 32; ++w) { const int c = sWarp[w]; before += w < warp ? c : 0; cnt += c; }
        if (staged + cnt > TK_CAP) {   // no room for this chunk: fold what is staged into the running best first
            bitonic_sort_2048(sc, si);
            sc[TK_CAP + tid] = ~0ull; si[TK_CAP + tid] = ~0ull;
            staged = 0;
        }
        __syncthreads();
        if (valid) {
            const int at = TK_CAP + staged + before + __popc(m & ((1u << lane) - 1u));
            sc[at] = inCost[i]; si[at] = (uint64_t)(firstCandidate + i);
        }
        staged += cnt; totalValid += cnt;
        __syncthreads();
    }
    if (staged > 0) bitonic_sort_2048(sc, si);
    __syncthreads();
    if (tid == 0 && totalValid && outValid) atomicAdd(outValid, totalValid);
    if (tid < k) { outCost[(int64_t)blockIdx.x * k + tid] = sc[tid]; outId[(int64_t)blockIdx.x * k + tid] = si[tid]; }
}

// Pass 2.  The per-CTA lists are sorted and all keys are distinct (the id is unique), so the global rank of an entry is
// its own index plus, for every other list, the number of entries that sort before it: L binary searches per entry,
// spread over the whole grid; entries of rank < k are scattered to their place.
__global__ void __launch_bounds__(256) topk_rank_merge(const uint64_t* partCost, const uint64_t* partId, int lists, int k,
                                                       uint64_t* outCost, uint64_t* outId) {
    const int64_t total = (int64_t)lists * k;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(e / k), j = (int)(e - (int64_t)r * k);
        const uint64_t c = partCost[e], id = partId[e];
        if (id == ~0ull) continue;  // padding
        int rank = j;
        for (int q = 0; q < lists && rank < k; ++q) {
            if (q == r) continue;
            const uint64_t* oc = partCost + (int64_t)q * k;
            const uint64_t* oi = partId + (int64_t)q * k;
            int lo = 0, hi = k;  // first index whose key is not < (c, id)
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (key_less(oc[mid], oi[mid], c, id)) lo = mid + 1; else hi = mid;
            }
            rank += lo;
        }
        if (rank < k) { outCost[rank] = c; outId[rank] = id; }
    }
}

__global__ void topk_fill(uint64_t* outCost, uint64_t* outId, int k) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < k) { outCost[i] = ~0ull; outId[i] = ~0ull; }
}

static int topk_ctas(int64_t nCand) {
    // few, long slices: valid keys are compacted before they reach the sorter, so a CTA usually sorts once however many
    // candidates it streams, and the rank merge costs (lists - 1) dependent binary searches per entry
    int64_t c = (nCand + 8 * TK_CAP - 1) / (8 * TK_CAP);
    if (c < 1) c = 1;
    if (c > 148) c = 148;
    return (int)c;
}

size_t topk_scratch_bytes(int64_t nCand, int k) { return (size_t)topk_ctas(nCand) * k * 2 * sizeof(uint64_t) + 16; }

cudaError_t launch_topk(const uint64_t* cost, const uint32_t* flags, int64_t nCand, int64_t firstCandidate, int k,
                        uint64_t* outKeyCost, uint64_t* outKeyId, int32_t* outCount, void* scratch, size_t scratchBytes,
                        cudaStream_t stream) {
    if (k < 1 || k > TK_CAP) return cudaErrorInvalidValue;
    const int ctas = topk_ctas(nCand);
    if (scratchBytes < topk_scratch_bytes(nCand, k)) return cudaErrorInvalidValue;
    uint64_t* partCost = reinterpret_cast<uint64_t*>(scratch);
    uint64_t* partId = partCost + (size_t)ctas * k;
    cudaMemsetAsync(outCount, 0, sizeof(int32_t), stream);
    topk_select<<<ctas, TK_THREADS, 0, stream>>>(cost, flags, nCand, firstCandidate, k, partCost, partId, outCount);
    topk_fill<<<(k + 255) / 256, 256, 0, stream>>>(outKeyCost, outKeyId, k);
    const int64_t total = (int64_t)ctas * k;
    topk_rank_merge<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(partCost, partId, ctas, k, outKeyCost, outKeyId);
    return cudaGetLastError();
}

}  // namespace bcu
