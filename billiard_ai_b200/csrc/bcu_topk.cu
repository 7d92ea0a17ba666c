// K2 topk: block-level top-k on (cost, candidate id) among VALID candidates.
//
// The reference returns solutions in ascending cost (process::Comparator, process.hpp:19-25, and
// SearchNode::_cost, type.hpp:572); ties are broken here by the global candidate id so that the
// selection is deterministic and identical on every rank.
//
// Valid strikes are a small share of a level, so the selection works on a compacted list:
//   topk_compact     every valid (cost, id) is appended to one global list (one atomicAdd per warp); V = their number.
//   topk_rank_small  V <= TK_LIST: the rank of a key is the number of keys that sort before it -- V * V comparisons spread
//                    over the grid (a warp per key), keys of rank < k scattered straight to their place.  No sort.
//   V > TK_LIST (dense levels): the streaming sorter below.  Pass 1 (topk_select): each CTA compacts the valid keys of
//   its slice into a 2048-entry bitonic sorter in shared memory, keeping a running best-1024.  Pass 2 (topk_rank_merge):
//   the sorted per-CTA lists are merged by rank.
// All four kernels are launched every time; the pair that does not apply reads V and returns (the host never syncs).
#include <cuda_runtime.h>

#include "bcu_kernels.cuh"

namespace bcu {

constexpr int TK_THREADS = 1024;
constexpr int TK_CAP = 1024;  // max k
constexpr int TK_LIST = 16384;  // compacted list: above this many valid keys the streaming sorter takes over
constexpr int TK_LANES = 32;    // lanes that share the comparisons of one key in topk_rank_small (a warp per key)

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

// Compaction of the valid keys (order within the list is irrelevant: ranks come from the keys).  *outValid counts every
// valid key, also those beyond the list's capacity.  The first k threads pre-fill the output with padding.
__global__ void __launch_bounds__(256) topk_compact(const uint64_t* __restrict__ inCost, const uint32_t* __restrict__ inFlags,
                                                    int64_t n, int64_t firstCandidate, int k, ulonglong2* __restrict__ list,
                                                    int32_t* outValid, uint64_t* outCost, uint64_t* outId) {
    const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t0 < k) { outCost[t0] = ~0ull; outId[t0] = ~0ull; }
    const int lane = threadIdx.x & 31;
    const int64_t step = (int64_t)gridDim.x * blockDim.x;
    const int64_t nRound = (n + 31) & ~(int64_t)31;   // whole warps stay in the loop together
    for (int64_t i = t0; i < nRound; i += step) {
        const bool valid = i < n && (inFlags[i] & F_VALID);
        const unsigned m = __ballot_sync(0xffffffffu, valid);
        if (m == 0) continue;
        int base = 0;
        if (lane == 0) base = atomicAdd(outValid, __popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (valid) {
            const int at = base + __popc(m & ((1u << lane) - 1u));
            if (at < TK_LIST) list[at] = make_ulonglong2(inCost[i], (uint64_t)(firstCandidate + i));
        }
    }
}

// Rank by counting.  Keys are distinct (the id is unique), so rank(key) = #{keys that sort before it}.  A warp shares
// one key and strides through the list together (32 consecutive entries per step, L1 / L2 resident); the kernel is
// bound by load latency, so it runs one key per warp with the SMs full of warps rather than fewer, longer threads.
__global__ void __launch_bounds__(256) topk_rank_small(const ulonglong2* __restrict__ list, const int32_t* __restrict__ nValid,
                                                       int k, uint64_t* outCost, uint64_t* outId) {
    const int V = *nValid;
    if (V > TK_LIST) return;                           // the streaming sorter handles this level
    const int e = (blockIdx.x * blockDim.x + threadIdx.x) / TK_LANES;
    const int q = threadIdx.x & (TK_LANES - 1);
    if ((int)(blockIdx.x * blockDim.x / TK_LANES) >= V) return;
    const bool live = e < V;
    const ulonglong2 me = live ? list[e] : make_ulonglong2(0ull, 0ull);
    const unsigned __int128 mine = ((unsigned __int128)me.x << 64) | me.y;
    int rank = 0;
    int jb = 0;   // warp-uniform base index: every lane makes the same trips (there are shuffles inside)
    // A key whose partial count reaches k is out whatever the rest of the list holds: the warp checks every 1024 entries
    // and leaves early (the exact rank only matters below k), which removes about half of the comparisons.
    while (jb + 4 * TK_LANES <= V) {
        #pragma unroll 1
        for (int it = 0; it < 8 && jb + 4 * TK_LANES <= V; ++it, jb += 4 * TK_LANES) {
            const int j = jb + q;
            const ulonglong2 a = list[j], b = list[j + TK_LANES], c = list[j + 2 * TK_LANES], d = list[j + 3 * TK_LANES];
            rank += ((((unsigned __int128)a.x << 64) | a.y) < mine) ? 1 : 0;
            rank += ((((unsigned __int128)b.x << 64) | b.y) < mine) ? 1 : 0;
            rank += ((((unsigned __int128)c.x << 64) | c.y) < mine) ? 1 : 0;
            rank += ((((unsigned __int128)d.x << 64) | d.y) < mine) ? 1 : 0;
        }
        int total = rank;
        #pragma unroll
        for (int s = 1; s < TK_LANES; s <<= 1) total += __shfl_xor_sync(0xffffffffu, total, s);
        if (total >= k) return;   // warp-uniform
    }
    for (int j = jb + q; j < V; j += TK_LANES) {
        const ulonglong2 a = list[j];
        rank += ((((unsigned __int128)a.x << 64) | a.y) < mine) ? 1 : 0;
    }
    #pragma unroll
    for (int s = 1; s < TK_LANES; s <<= 1) rank += __shfl_xor_sync(0xffffffffu, rank, s);
    if (live && q == 0 && rank < k) { outCost[rank] = me.x; outId[rank] = me.y; }
}

// Pass 1.  Each CTA streams its slice of the candidates: VALID keys are compacted (ballot + prefix) into the upper half
// of the 2048-entry sorter, and only when that half could overflow (or at the end) the sorter runs, leaving the running
// best-1024 in the lower half.  With the usual few-percent share of valid strikes a CTA sorts once.
__global__ void __launch_bounds__(TK_THREADS) topk_select(const uint64_t* inCost, const uint32_t* inFlags, int64_t n,
                                                          int64_t firstCandidate, int k, uint64_t* outCost, uint64_t* outId,
                                                          const int32_t* nValid) {
    if (*nValid <= TK_LIST) return;                    // topk_rank_small has done this level
    __shared__ uint64_t sc[2 * TK_CAP];
    __shared__ uint64_t si[2 * TK_CAP];
    __shared__ int sWarp[TK_THREADS / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    sc[tid] = ~0ull; si[tid] = ~0ull;
    sc[TK_CAP + tid] = ~0ull; si[TK_CAP + tid] = ~0ull;
    const int64_t per = (n + gridDim.x - 1) / gridDim.x;
    const int64_t lo = per * blockIdx.x;
    const int64_t hi = lo + per < n ? lo + per : n;
    int staged = 0;   // block-uniform
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
        staged += cnt;
        __syncthreads();
    }
    if (staged > 0) bitonic_sort_2048(sc, si);
    __syncthreads();
    if (tid < k) { outCost[(int64_t)blockIdx.x * k + tid] = sc[tid]; outId[(int64_t)blockIdx.x * k + tid] = si[tid]; }
}

// Pass 2.  The per-CTA lists are sorted and all keys are distinct (the id is unique), so the global rank of an entry is
// its own index plus, for every other list, the number of entries that sort before it: L binary searches per entry,
// spread over the whole grid; entries of rank < k are scattered to their place.
__global__ void __launch_bounds__(256) topk_rank_merge(const uint64_t* partCost, const uint64_t* partId, int lists, int k,
                                                       uint64_t* outCost, uint64_t* outId, const int32_t* nValid) {
    if (*nValid <= TK_LIST) return;
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

static int topk_ctas(int64_t nCand) {
    // few, long slices: valid keys are compacted before they reach the sorter, so a CTA usually sorts once however many
    // candidates it streams, and the rank merge costs (lists - 1) dependent binary searches per entry
    int64_t c = (nCand + 8 * TK_CAP - 1) / (8 * TK_CAP);
    if (c < 1) c = 1;
    if (c > 148) c = 148;
    return (int)c;
}

size_t topk_scratch_bytes(int64_t nCand, int k) {
    return (size_t)TK_LIST * sizeof(ulonglong2) + (size_t)topk_ctas(nCand) * k * 2 * sizeof(uint64_t) + 16;
}

cudaError_t launch_topk(const uint64_t* cost, const uint32_t* flags, int64_t nCand, int64_t firstCandidate, int k,
                        uint64_t* outKeyCost, uint64_t* outKeyId, int32_t* outCount, void* scratch, size_t scratchBytes,
                        cudaStream_t stream) {
    if (k < 1 || k > TK_CAP) return cudaErrorInvalidValue;
    const int ctas = topk_ctas(nCand);
    if (scratchBytes < topk_scratch_bytes(nCand, k)) return cudaErrorInvalidValue;
    ulonglong2* list = reinterpret_cast<ulonglong2*>(scratch);
    uint64_t* partCost = reinterpret_cast<uint64_t*>(list + TK_LIST);
    uint64_t* partId = partCost + (size_t)ctas * k;
    cudaMemsetAsync(outCount, 0, sizeof(int32_t), stream);
    int64_t cc = (nCand + 1023) / 1024;   // 4 candidates per thread at least, one wave at most
    if (cc < (k + 255) / 256) cc = (k + 255) / 256;
    if (cc > 148 * 4) cc = 148 * 4;
    topk_compact<<<(unsigned)cc, 256, 0, stream>>>(cost, flags, nCand, firstCandidate, k, list, outCount, outKeyCost, outKeyId);
    topk_rank_small<<<TK_LIST * TK_LANES / 256, 256, 0, stream>>>(list, outCount, k, outKeyCost, outKeyId);
    topk_select<<<ctas, TK_THREADS, 0, stream>>>(cost, flags, nCand, firstCandidate, k, partCost, partId, outCount);
    const int64_t total = (int64_t)ctas * k;
    topk_rank_merge<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(partCost, partId, ctas, k, outKeyCost, outKeyId, outCount);
    return cudaGetLastError();
}

}  // namespace bcu

// ---------------------------------------------------------------------------------------------------------------------
// bcu_topk_static: the same selection for callers that cannot allocate (code under CUDA graph capture, serving loops and
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
//
// The scratch of launch_topk -- the compacted list of valid keys and the per-CTA sorted lists of the streaming sorter --
// lives in ONE file-scope device array per GPU that loads the library (held until exit): TK_LIST keys of the list plus
// 148 lists of up to 1024 keys, 2 687 232 bytes.  Nothing in it has to start at zero: the list is written before it is read
// up to the valid count (which lives in the caller's d_out_count, cleared here with a memset node), and every per-CTA list
// is written whole by the CTA that owns it before the merge reads it.  Every call on a device shares the array, so two calls
// that overlap on the GPU (different streams, no dependency) race; calls on one stream are ordered and fine.
// ---------------------------------------------------------------------------------------------------------------------
#include <cstdio>
#include <mutex>

#include "../../include/billiard_cuda.h"

namespace bcu {
constexpr size_t TOPK_STATIC_BYTES = (size_t)TK_LIST * sizeof(ulonglong2) + (size_t)148 * TK_CAP * 2 * sizeof(uint64_t) + 256;
static __device__ __align__(256) unsigned char g_topkStaticScratch[TOPK_STATIC_BYTES];

constexpr int TOPK_STATIC_MAX_DEVICES = 64;
static std::mutex g_topkStaticMutex;
static void* g_topkStaticPtr[TOPK_STATIC_MAX_DEVICES] = {};   // device address of the array on device i (0 = not asked yet)

// In host code `g_topkStaticScratch` names a host-side shadow of the symbol, not device memory: the address a kernel may
// use comes from cudaGetSymbolAddress, asked once per device with that device current.
static cudaError_t topk_static_scratch(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= TOPK_STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_topkStaticMutex);
    if (!g_topkStaticPtr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_topkStaticScratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        g_topkStaticPtr[dev] = p;
    }
    *out = g_topkStaticPtr[dev];
    return cudaSuccess;
}
}  // namespace bcu

extern "C" int bcu_topk_static(const uint64_t* d_cost, const uint32_t* d_flags, int64_t n, int64_t first_candidate, int32_t k,
                               uint64_t* d_out_cost, uint64_t* d_out_id, int32_t* d_out_count, void* stream,
                               char* err, int32_t err_cap) {
    using namespace bcu;
    auto fail = [&](int code, const char* what, const char* detail) {
        if (err && err_cap > 0) snprintf(err, (size_t)err_cap, "bcu_topk_static: %s%s%s", what, detail ? ": " : "", detail ? detail : "");
        return code;
    };
    if (err && err_cap > 0) err[0] = 0;
    if (!d_cost || !d_flags || !d_out_cost || !d_out_id || !d_out_count) return fail(BCU_ERR_INVALID, "null pointer argument", nullptr);
    if (n < 0 || n > BCU_TOPK_STATIC_MAX_CANDIDATES) {
        char buf[160];
        snprintf(buf, sizeof buf, "n = %lld candidates, the static scratch serves 0 .. %lld (use bcu_topk beyond that)", (long long)n,
                 (long long)BCU_TOPK_STATIC_MAX_CANDIDATES);
        return fail(BCU_ERR_INVALID, "size beyond the maximum", buf);
    }
    if (k < 1 || k > BCU_TOPK_STATIC_MAX_K) {
        char buf[96];
        snprintf(buf, sizeof buf, "k = %d, allowed 1 .. %d", (int)k, (int)BCU_TOPK_STATIC_MAX_K);
        return fail(BCU_ERR_INVALID, "size beyond the maximum", buf);
    }
    static_assert(BCU_TOPK_STATIC_MAX_K <= TK_CAP, "k limit");
    if (topk_scratch_bytes(n, k) > TOPK_STATIC_BYTES) return fail(BCU_ERR_INVALID, "size beyond the maximum", "scratch");
    void* scratch = nullptr;
    cudaError_t ce = topk_static_scratch(&scratch);
    if (ce != cudaSuccess) {
        const int code = (ce == cudaErrorNoDevice || ce == cudaErrorInsufficientDriver) ? BCU_ERR_NO_DEVICE : BCU_ERR_CUDA;
        return fail(code, "cudaGetSymbolAddress", cudaGetErrorString(ce));
    }
    ce = launch_topk(d_cost, d_flags, n, first_candidate, k, d_out_cost, d_out_id, d_out_count, scratch, TOPK_STATIC_BYTES,
                     (cudaStream_t)stream);
    if (ce != cudaSuccess) return fail(BCU_ERR_CUDA, "launch", cudaGetErrorString(ce));
    return BCU_OK;
}
