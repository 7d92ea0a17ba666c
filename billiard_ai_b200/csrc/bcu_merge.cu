// Beam records and their multi-GPU merge.
//
//   pack_records : the local top-k (cost, id) list of one GPU + the end state of each selected candidate
//                  -> one contiguous record block  [cost u64 x k | id u64 x k | potted u32 x k | time f32 x k |
//                                                   pos f32 x k x NB x 2]   (entries past the valid count: cost = id = ~0)
//   merge_records: G such blocks (after ncclAllGather) -> the global top-k block, same layout, ascending (cost, id).
//                  Every rank runs it on identical input, so every rank holds the identical beam (no broadcast).
//
// The order is the reference's solution order: ascending SearchNode::_cost (process::Comparator,
// process.hpp:19-25), ties broken by the global candidate id for determinism.
#include <cuda_runtime.h>

#include "bcu_kernels.cuh"

namespace bcu {

size_t record_block_bytes(int k, int nBalls) {
    return (size_t)k * (8 + 8 + 4 + 4 + (size_t)nBalls * 2 * 4);
}

struct BlockView {
    uint64_t* cost;
    uint64_t* id;
    uint32_t* potted;
    float* time;
    float* pos;
};
__host__ __device__ inline BlockView view_block(void* base, int k) {
    BlockView v;
    char* p = (char*)base;
    v.cost = (uint64_t*)p; p += (size_t)k * 8;
    v.id = (uint64_t*)p; p += (size_t)k * 8;
    v.potted = (uint32_t*)p; p += (size_t)k * 4;
    v.time = (float*)p; p += (size_t)k * 4;
    v.pos = (float*)p;
    return v;
}

__global__ void pack_records_kernel(const uint64_t* keyCost, const uint64_t* keyId, int k, int64_t firstCandidate,
                                    const uint32_t* potted, const float* finalTime, const float* finalPos, int NB,
                                    void* out) {
    BlockView v = view_block(out, k);
    const int j = blockIdx.x;
    const uint64_t id = keyId[j];
    const bool valid = id != ~0ull;
    const int64_t local = valid ? (int64_t)id - firstCandidate : 0;
    if (threadIdx.x == 0) {
        v.cost[j] = keyCost[j];
        v.id[j] = id;
        v.potted[j] = valid ? potted[local] : 0u;
        v.time[j] = valid ? finalTime[local] : 0.0f;
    }
    for (int i = threadIdx.x; i < NB * 2; i += blockDim.x)
        v.pos[(size_t)j * NB * 2 + i] = valid ? finalPos[(size_t)local * NB * 2 + i] : 0.0f;
}

cudaError_t launch_pack_records(const uint64_t* keyCost, const uint64_t* keyId, int k, int64_t firstCandidate,
                                const uint32_t* potted, const float* finalTime, const float* finalPos, int nBalls,
                                void* out, cudaStream_t stream) {
    pack_records_kernel<<<k, 64, 0, stream>>>(keyCost, keyId, k, firstCandidate, potted, finalTime, finalPos, nBalls, out);
    return cudaGetLastError();
}

// One warp per gathered entry: rank r's list is already sorted, so the global rank of entry (r, j) is
//   j + sum over r' != r of #{ entries of r' that sort before it }  -- G - 1 binary searches (every lane runs the same
// search: no divergence), then the warp copies the record to its place.
__global__ void __launch_bounds__(256) merge_records_kernel(const void* gathered, int G, int k, int NB, size_t blockBytes,
                                                            void* out) {
    BlockView o = view_block(out, k);
    const int lane = threadIdx.x & 31;
    const int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (e >= G * k) return;
    const int r = e / k, j = e - r * k;
    BlockView mine = view_block((char*)gathered + (size_t)r * blockBytes, k);
    const uint64_t c = mine.cost[j], id = mine.id[j];
    if (id == ~0ull) return;  // padding
    int rank = j;
    for (int q = 0; q < G && rank < k; ++q) {
        if (q == r) continue;
        BlockView other = view_block((char*)gathered + (size_t)q * blockBytes, k);
        int lo = 0, hi = k;  // first index whose key is not < (c, id)
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            const uint64_t oc = other.cost[mid], oi = other.id[mid];
            const bool less = oc < c || (oc == c && oi < id);
            if (less) lo = mid + 1; else hi = mid;
        }
        rank += lo;
    }
    if (rank >= k) return;
    if (lane == 0) {
        o.cost[rank] = c;
        o.id[rank] = id;
        o.potted[rank] = mine.potted[j];
        o.time[rank] = mine.time[j];
    }
    for (int i = lane; i < NB * 2; i += 32) o.pos[(size_t)rank * NB * 2 + i] = mine.pos[(size_t)j * NB * 2 + i];
}

// padding for the slots the merge leaves untouched (fewer than k valid entries over all ranks)
__global__ void __launch_bounds__(256) fill_block_kernel(void* out, int k, int NB) {
    BlockView o = view_block(out, k);
    const int t = threadIdx.x + blockIdx.x * blockDim.x, step = blockDim.x * gridDim.x;
    for (int j = t; j < k; j += step) { o.cost[j] = ~0ull; o.id[j] = ~0ull; o.potted[j] = 0u; o.time[j] = 0.0f; }
    for (int i = t; i < k * NB * 2; i += step) o.pos[i] = 0.0f;
}

cudaError_t launch_merge_records(const void* gathered, int G, int k, int nBalls, void* out, cudaStream_t stream) {
    fill_block_kernel<<<(k * nBalls * 2 + 255) / 256, 256, 0, stream>>>(out, k, nBalls);
    merge_records_kernel<<<(G * k + 7) / 8, 256, 0, stream>>>(gathered, G, k, nBalls, record_block_bytes(k, nBalls), out);
    return cudaGetLastError();
}

// ---- measurement helpers ----------------------------------------------------------------------------------
// Issue-rate microbenchmarks for the roofline denominators (MEASURED_PEAKS.json has no FP64 / FP32 vector peaks):
//   kind 0: DADD+DMUL (un-fused, what this path may use)   kind 1: DFMA   kind 2: FADD+FMUL   kind 3: FFMA
template <int KIND>
__global__ void __launch_bounds__(256) issue_rate_kernel(int iters, double* sink) {
    if (KIND < 2) {
        double a0 = threadIdx.x * 1e-9 + 1.0, a1 = a0 + 0.1, a2 = a0 + 0.2, a3 = a0 + 0.3;
        double a4 = a0 + 0.4, a5 = a0 + 0.5, a6 = a0 + 0.6, a7 = a0 + 0.7;
        const double m = 0.999999, c = 1e-7;
        for (int i = 0; i < iters; ++i) {
            if (KIND == 0) {
                a0 = __dadd_rn(__dmul_rn(a0, m), c); a1 = __dadd_rn(__dmul_rn(a1, m), c);
                a2 = __dadd_rn(__dmul_rn(a2, m), c); a3 = __dadd_rn(__dmul_rn(a3, m), c);
                a4 = __dadd_rn(__dmul_rn(a4, m), c); a5 = __dadd_rn(__dmul_rn(a5, m), c);
                a6 = __dadd_rn(__dmul_rn(a6, m), c); a7 = __dadd_rn(__dmul_rn(a7, m), c);
            } else {
                a0 = __fma_rn(a0, m, c); a1 = __fma_rn(a1, m, c); a2 = __fma_rn(a2, m, c); a3 = __fma_rn(a3, m, c);
                a4 = __fma_rn(a4, m, c); a5 = __fma_rn(a5, m, c); a6 = __fma_rn(a6, m, c); a7 = __fma_rn(a7, m, c);
            }
        }
        if (a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 == 12345.678) *sink = a0;
    } else {
        float a0 = threadIdx.x * 1e-6f + 1.0f, a1 = a0 + 0.1f, a2 = a0 + 0.2f, a3 = a0 + 0.3f;
        float a4 = a0 + 0.4f, a5 = a0 + 0.5f, a6 = a0 + 0.6f, a7 = a0 + 0.7f;
        const float m = 0.99999f, c = 1e-6f;
        for (int i = 0; i < iters; ++i) {
            if (KIND == 2) {
                a0 = __fadd_rn(__fmul_rn(a0, m), c); a1 = __fadd_rn(__fmul_rn(a1, m), c);
                a2 = __fadd_rn(__fmul_rn(a2, m), c); a3 = __fadd_rn(__fmul_rn(a3, m), c);
                a4 = __fadd_rn(__fmul_rn(a4, m), c); a5 = __fadd_rn(__fmul_rn(a5, m), c);
                a6 = __fadd_rn(__fmul_rn(a6, m), c); a7 = __fadd_rn(__fmul_rn(a7, m), c);
            } else {
                a0 = __fmaf_rn(a0, m, c); a1 = __fmaf_rn(a1, m, c); a2 = __fmaf_rn(a2, m, c); a3 = __fmaf_rn(a3, m, c);
                a4 = __fmaf_rn(a4, m, c); a5 = __fmaf_rn(a5, m, c); a6 = __fmaf_rn(a6, m, c); a7 = __fmaf_rn(a7, m, c);
            }
        }
        if (a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 == 12345.678f) *sink = a0;
    }
}

// returns the number of arithmetic INSTRUCTIONS per thread-iteration (8 chains x (2 | 1))
int launch_issue_rate(int kind, int ctas, int iters, double* sink, cudaStream_t stream) {
    switch (kind) {
        case 0: issue_rate_kernel<0><<<ctas, 256, 0, stream>>>(iters, sink); return 16;
        case 1: issue_rate_kernel<1><<<ctas, 256, 0, stream>>>(iters, sink); return 8;
        case 2: issue_rate_kernel<2><<<ctas, 256, 0, stream>>>(iters, sink); return 16;
        default: issue_rate_kernel<3><<<ctas, 256, 0, stream>>>(iters, sink); return 8;
    }
}

__global__ void flush_kernel(float4* buf, size_t n, float v) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        buf[i] = make_float4(v, v, v, v);
}
cudaError_t launch_flush(void* buf, size_t bytes, cudaStream_t stream) {
    flush_kernel<<<148 * 8, 256, 0, stream>>>((float4*)buf, bytes / 16, 1.0f);
    return cudaGetLastError();
}

}  // namespace bcu
