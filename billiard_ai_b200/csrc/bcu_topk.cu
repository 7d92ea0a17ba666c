// K2 topk: block-level top-k on (cost, candidate id) among VALID candidates.
//
// The reference returns solutions in ascending cost (process::Comparator, process.hpp:19-25, and
// SearchNode::_cost, type.hpp:572); ties are broken here by the global candidate id so that the
// selection is deterministic and identical on every rank.
//
// Pass 1: each CTA streams its slice of the candidates through a 2048-entry bitonic sorter in shared
// memory, keeping a running best-1024.  Pass 2: one CTA merges the per-CTA lists the same way.
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

// MODE 0: input = (cost[], flags[]) of a simulate launch, id = firstCandidate + index, only VALID entries
// MODE 1: input = (keyCost[], keyId[]) lists from pass 1 (invalid entries are ~0)
template <int MODE>
__global__ void __launch_bounds__(TK_THREADS) topk_pass(const uint64_t* inCost, const uint32_t* inFlags,
                                                        const uint64_t* inId, int64_t n, int64_t firstCandidate, int k,
                                                        uint64_t* outCost, uint64_t* outId, int32_t* outValid) {
    __shared__ uint64_t sc[2 * TK_CAP];
    __shared__ uint64_t si[2 * TK_CAP];
    __shared__ int sValid;
    const int tid = threadIdx.x;
    if (tid == 0) sValid = 0;
    sc[tid] = ~0ull; si[tid] = ~0ull;
    const int64_t per = (n + gridDim.x - 1) / gridDim.x;
    const int64_t lo = per * blockIdx.x;
    const int64_t hi = lo + per < n ? lo + per : n;
    int myValid = 0;
    for (int64_t base = lo; base < hi; base += TK_CAP) {
        const int64_t i = base + tid;
        uint64_t c = ~0ull, id = ~0ull;
        if (i < hi) {
            if (MODE == 0) {
                if (inFlags[i] & F_VALID) { c = inCost[i]; id = (uint64_t)(firstCandidate + i); myValid++; }
            } else {
                c = inCost[i]; id = inId[i];
            }
        }
        sc[TK_CAP + tid] = c; si[TK_CAP + tid] = id;
        bitonic_sort_2048(sc, si);  // best TK_CAP now in the first half (second half is overwritten next round)
    }
    if (MODE == 0 && outValid) {
        __syncthreads();
        if (myValid) atomicAdd(&sValid, myValid);
        __syncthreads();
        if (tid == 0 && sValid) atomicAdd(outValid, sValid);
    }
    __syncthreads();
    if (tid < k) { outCost[(int64_t)blockIdx.x * k + tid] = sc[tid]; outId[(int64_t)blockIdx.x * k + tid] = si[tid]; }
}

static int topk_ctas(int64_t nCand) {
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
    topk_pass<0><<<ctas, TK_THREADS, 0, stream>>>(cost, flags, nullptr, nCand, firstCandidate, k, partCost, partId, outCount);
    topk_pass<1><<<1, TK_THREADS, 0, stream>>>(partCost, nullptr, partId, (int64_t)ctas * k, 0, k, outKeyCost, outKeyId, nullptr);
    return cudaGetLastError();
}

}  // namespace bcu
