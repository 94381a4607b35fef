// Two more shapes of the Filter answer (K3), for callers that do not want the whole [P][N] bitmap:
//
//   compact  : the predicate  key[n] >= req[p]  (NodeInfo.Assume, pkg/cache/nodeinfo.go:113-137, with
//              key[n] = max available memory over the node's devices) is monotone in req, so the whole
//              [P][N] answer is carried by one number per node and one per pod,
//                  rank[n]   = #{ q : req[q] <= key[n] }        (0 for a node that can never pass)
//                  cutoff[p] = 1 + #{ q : req[q] < req[p] }
//              with   pass(p, n)  <=>  rank[n] >= cutoff[p]   for every 64-bit request:
//                key[n] >= req[p]: every q counted by cutoff[p] - 1, and p itself, has req <= key[n];
//                key[n] <  req[p]: every q counted by rank[n] has req[q] < req[p].
//              2 bytes per node + 2 per pod cross PCIe instead of a bit per (pod, node).
//   subset   : the candidate list of one Filter request (`for _, nodeName := range *args.NodeNames`,
//              pkg/scheduler/predicate.go:21): bit c of row p answers node subset[c].
#include "common.cuh"
#include "kernels.h"
#include "state_device.cuh"

namespace gs {

using u64 = unsigned long long;

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------
// compact, step 1: part[s][p] = #{ q in column slice s : req[q] < req[p] }.  All pairs, spread over
// (row blocks x column slices) CTAs; a slice is staged in shared memory and every thread of the block
// reads the same element at the same time (broadcast).  Requests below 2^32 -- all of them unless a
// pod's limits sum to a negative number (pkg/cache/nodeinfo.go:217) -- compare as 32-bit values.
// ------------------------------------------------------------------------------------------------
constexpr int kCountThreads = 256;
constexpr uint32_t kCountChunk = 1024;  // columns staged at a time

__global__ void __launch_bounds__(kCountThreads)
k_req_count(const u64* __restrict__ req, uint32_t n_pods, uint32_t slice, uint32_t* __restrict__ part) {
    __shared__ __align__(16) u64 col64[kCountChunk];
    __shared__ __align__(16) uint32_t col32[kCountChunk];
    const uint32_t tid = threadIdx.x;
    const uint32_t p = blockIdx.x * kCountThreads + tid;
    const uint32_t c_begin = blockIdx.y * slice;
    const uint32_t c_end = min(n_pods, c_begin + slice);
    pdl_launch_dependents();
    const u64 mine = p < n_pods ? req[p] : 0ull;
    uint32_t cnt = 0;
    for (uint32_t c0 = c_begin; c0 < c_end; c0 += kCountChunk) {
        const uint32_t len = min(kCountChunk, c_end - c0);
        const uint32_t len4 = (len + 3u) & ~3u;
        int big = (mine >> 32) != 0;
        for (uint32_t i = tid; i < len4; i += kCountThreads) {
            const u64 v = i < len ? req[c0 + i] : ~0ull;  // padding is never below anything
            col64[i] = v;
            col32[i] = (v >> 32) ? 0xffffffffu : (uint32_t)v;
            big |= (i < len && (v >> 32) != 0);
        }
        const bool wide = __syncthreads_or(big) != 0;
        if (!wide) {
            const uint32_t m = (uint32_t)mine;
            const uint4* c4 = reinterpret_cast<const uint4*>(col32);
#pragma unroll 4
            for (uint32_t i = 0; i < len4 / 4; ++i) {
                const uint4 v = c4[i];
                cnt += (v.x < m) + (v.y < m) + (v.z < m) + (v.w < m);
            }
        } else {
            const ulonglong2* c2 = reinterpret_cast<const ulonglong2*>(col64);
#pragma unroll 4
            for (uint32_t i = 0; i < len4 / 2; ++i) {
                const ulonglong2 v = c2[i];
                cnt += (v.x < mine) + (v.y < mine);
            }
        }
        __syncthreads();  // the next chunk overwrites the stage
    }
    if (p < n_pods) part[(size_t)blockIdx.y * n_pods + p] = cnt;
}

// compact, step 2: lt = sum of the partial counts; cutoff[p] = lt + 1; the request goes to position
// lt + (arrival number among the pods with the same request) of the sorted array.  Equal requests have
// equal lt and distinct requests distinct lt, so the duplicate counter of a value lives at dup[lt]; the
// lanes of a warp that share a value go to it as one atomic.
__global__ void __launch_bounds__(256)
k_req_place(const u64* __restrict__ req, uint32_t n_pods, uint32_t splits, const uint32_t* __restrict__ part,
            uint32_t* __restrict__ dup, u64* __restrict__ sorted, uint16_t* __restrict__ cutoff) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31;
    pdl_launch_dependents();
    const u64 mine = p < n_pods ? req[p] : 0ull;
    pdl_wait();  // the partial counts
    uint32_t lt = 0xffffff00u + lane;  // idle lanes: a group of their own
    if (p < n_pods) {
        lt = 0;
        for (uint32_t s = 0; s < splits; ++s) lt += part[(size_t)s * n_pods + p];
    }
    const uint32_t same = __match_any_sync(0xffffffffu, lt);
    const uint32_t leader = __ffs(same) - 1;
    uint32_t first = 0;
    if (lane == leader && p < n_pods) first = atomicAdd(&dup[lt], (uint32_t)__popc(same));
    first = __shfl_sync(0xffffffffu, first, leader);
    if (p < n_pods) {
        sorted[lt + first + __popc(same & ((1u << lane) - 1u))] = mine;
        cutoff[p] = (uint16_t)(lt + 1u);
    }
}

// compact, step 3: one thread per node: key from the node's used[] row, rank = #{sorted requests <= key}
// by binary search (the 10k sorted requests stay in L1/L2).  The threads also re-arm the duplicate
// counters for the next call.
__global__ void __launch_bounds__(256)
k_node_rank(const u64* __restrict__ sorted, uint32_t n_pods, uint32_t* __restrict__ dup, const u64* __restrict__ used,
            const u64* __restrict__ cap, const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes, uint32_t dev_stride,
            uint16_t* __restrict__ rank) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    pdl_launch_dependents();
    bool eligible = false;
    const u64 key = node_key(used, cap, nodeinfo, n_nodes, dev_stride, i, &eligible);  // independent of the request side
    pdl_wait();  // the sorted requests
    if (i < n_pods) dup[i] = 0;
    if (i >= n_nodes) return;
    uint32_t lo = 0;
    if (eligible) {
        uint32_t len = n_pods;
        while (len > 0) {
            const uint32_t half = len >> 1;
            if (__ldg(sorted + lo + half) <= key) {
                lo += half + 1;
                len -= half + 1;
            } else {
                len = half;
            }
        }
    }
    rank[i] = (uint16_t)lo;
}

// ------------------------------------------------------------------------------------------------
// subset: a block owns 256 candidates (8 words of every row); each thread computes its candidate's key once
// and the block then walks the pods: one ballot per warp and pod.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_filter_subset(const u64* __restrict__ req, uint32_t n_pods, const uint32_t* __restrict__ subset, uint32_t n_cand,
                const u64* __restrict__ used, const u64* __restrict__ cap, const uint32_t* __restrict__ nodeinfo,
                uint32_t n_nodes, uint32_t dev_stride, uint32_t* __restrict__ out, uint32_t row_stride) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t word = c >> 5, words = (n_cand + 31u) >> 5;
    bool eligible = false;
    u64 key = 0;
    if (c < n_cand) key = node_key(used, cap, nodeinfo, n_nodes, dev_stride, subset[c], &eligible);
    for (uint32_t p = blockIdx.y; p < n_pods; p += gridDim.y) {
        const u64 r = __ldg(req + p);
        const uint32_t bits = __ballot_sync(0xffffffffu, eligible && key >= r);
        if (lane == 0 && word < words) out[(size_t)p * row_stride + word] = bits;
    }
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
uint32_t compact_splits(uint32_t n_pods, int sm_count) {
    const uint32_t row_blocks = (n_pods + kCountThreads - 1) / kCountThreads;
    if (row_blocks == 0) return 1;
    uint32_t want = (4u * (uint32_t)(sm_count > 0 ? sm_count : 1) + row_blocks - 1) / row_blocks;
    if (want < 1) want = 1;
    if (want > kCompactMaxSplits) want = kCompactMaxSplits;
    if (want > n_pods) want = n_pods;
    return want;
}

template <class... Params, class... Args>
static cudaError_t launch_pdl(void (*kernel)(Params...), dim3 grid, dim3 block, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // the launch itself overlaps the predecessor's tail
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<Params>(args)...);
}

cudaError_t launch_compact(const u64* d_req, uint32_t n_pods, uint32_t* d_part, u64* d_sorted, uint32_t* d_dup,
                           const u64* d_used, const u64* d_cap, const uint32_t* d_nodeinfo, uint32_t n_nodes,
                           uint32_t dev_stride, uint16_t* d_rank, uint16_t* d_cutoff, int sm_count, cudaStream_t st,
                           uint32_t* n_launches) {
    *n_launches = 0;
    if (n_pods == 0 || n_pods > GS_MAX_BATCH) return cudaErrorInvalidValue;
    const uint32_t splits = compact_splits(n_pods, sm_count);
    const uint32_t slice = ((n_pods + splits - 1) / splits + 3u) & ~3u;
    const uint32_t row_blocks = (n_pods + kCountThreads - 1) / kCountThreads;
    // the first kernel of the chain is an ordinary launch: it waits for everything before it in the stream
    k_req_count<<<dim3(row_blocks, (n_pods + slice - 1) / slice), kCountThreads, 0, st>>>(d_req, n_pods, slice, d_part);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    ++*n_launches;
    e = launch_pdl(k_req_place, dim3((n_pods + 255) / 256), dim3(256), st, d_req, n_pods, (n_pods + slice - 1) / slice,
                   (const uint32_t*)d_part, d_dup, d_sorted, d_cutoff);
    if (e != cudaSuccess) return e;
    ++*n_launches;
    const uint32_t n_threads = n_nodes > n_pods ? n_nodes : n_pods;
    e = launch_pdl(k_node_rank, dim3((n_threads + 255) / 256), dim3(256), st, (const u64*)d_sorted, n_pods, d_dup, d_used,
                   d_cap, d_nodeinfo, n_nodes, dev_stride, d_rank);
    if (e != cudaSuccess) return e;
    ++*n_launches;
    return cudaSuccess;
}

cudaError_t launch_filter_subset(const u64* d_req, uint32_t n_pods, const uint32_t* d_subset, uint32_t n_cand,
                                 const u64* d_used, const u64* d_cap, const uint32_t* d_nodeinfo, uint32_t n_nodes,
                                 uint32_t dev_stride, uint32_t* d_out, uint32_t row_stride, cudaStream_t st) {
    if (n_pods == 0 || n_cand == 0) return cudaErrorInvalidValue;
    const uint32_t xb = (n_cand + 255) / 256;
    uint32_t yb = (4u * 148u + xb - 1) / xb;  // enough blocks to fill the GPU when the candidate list is short
    if (yb > n_pods) yb = n_pods;
    if (yb > 65535u) yb = 65535u;
    if (yb < 1) yb = 1;
    k_filter_subset<<<dim3(xb, yb), 256, 0, st>>>(d_req, n_pods, d_subset, n_cand, d_used, d_cap, d_nodeinfo, n_nodes,
                                                  dev_stride, d_out, row_stride);
    return cudaGetLastError();
}

}  // namespace gs
