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
// compact, step 1: the requests in slices of kSlice, each slice sorted by one CTA (bitonic network in shared
// memory; the tail of the last slice is padded with all ones, which sorts behind every request).
// ------------------------------------------------------------------------------------------------
constexpr uint32_t kSlice = 1024;
constexpr uint32_t kMaxSlices = (GS_MAX_BATCH + kSlice - 1) / kSlice;

// one element per thread; exchanges at distances below 32 are shuffles, the fifteen at 32 and above go through
// shared memory
__global__ void __launch_bounds__(kSlice)
k_slice_sort(const u64* __restrict__ req, uint32_t n_pods, u64* __restrict__ slices, u64* __restrict__ req_copy) {
    __shared__ u64 v[kSlice];
    const uint32_t tid = threadIdx.x;
    const uint32_t base = blockIdx.x * kSlice;
    pdl_launch_dependents();
    u64 mine = base + tid < n_pods ? req[base + tid] : ~0ull;
    // requests that live in host memory are read over PCIe once: the next kernel takes them from this copy
    if (req_copy && base + tid < n_pods) req_copy[base + tid] = mine;
    for (uint32_t k = 2; k <= kSlice; k <<= 1) {
        const bool asc = (tid & k) == 0;
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            u64 other;
            if (j >= 32) {
                __syncthreads();  // everybody has read what the previous exchange left
                v[tid] = mine;
                __syncthreads();
                other = v[tid ^ j];
            } else {
                other = __shfl_xor_sync(0xffffffffu, mine, j);
            }
            const bool lower = (tid & j) == 0;
            // the lower index keeps the smaller value in an ascending block, the larger in a descending one
            const bool take_min = lower == asc;
            mine = take_min ? (other < mine ? other : mine) : (other > mine ? other : mine);
        }
    }
    slices[base + tid] = mine;
}

// compact, step 2: lt[p] = #{q : req[q] < req[p]} = the sum over the sorted slices of a lower bound (the searches
// of a thread are independent chains of eleven loads from L1/L2); cutoff[p] = lt + 1; the request goes to position
// lt + (arrival number among the pods with the same request) of the fully sorted array.  Equal requests have equal
// lt and distinct requests distinct lt, so the duplicate counter of a value lives at dup[lt]; the lanes of a warp
// that share a value go to it as one atomic.
__global__ void __launch_bounds__(64)
k_req_place(const u64* __restrict__ req_early, const u64* __restrict__ req_late, uint32_t n_pods, uint32_t n_slices,
            const u64* __restrict__ g_slices, uint32_t* __restrict__ dup, u64* __restrict__ sorted,
            uint16_t* __restrict__ cutoff) {
    extern __shared__ __align__(16) u64 slices[];  // every block keeps all the slices at hand: the searches are dependent loads
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t lane = threadIdx.x & 31;
    pdl_launch_dependents();
    // the request: from the caller's device array before the wait, or after it from the copy k_slice_sort made
    u64 mine = (req_early && p < n_pods) ? req_early[p] : 0ull;
    pdl_wait();  // the sorted slices
    if (!req_early && p < n_pods) mine = req_late[p];
    {
        // one bulk asynchronous copy (TMA) per slice, all in flight together, completion on an mbarrier
        __shared__ __align__(8) uint64_t bar;
        const uint32_t bar_addr = (uint32_t)__cvta_generic_to_shared(&bar);
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_addr) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_addr), "r"(n_slices * kSlice * 8u) : "memory");
            for (uint32_t sl = 0; sl < n_slices; ++sl)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                 (uint32_t)__cvta_generic_to_shared(slices + sl * kSlice)),
                             "l"(g_slices + sl * kSlice), "r"(kSlice * 8u), "r"(bar_addr)
                             : "memory");
        }
        __syncthreads();  // the barrier is initialised before anybody waits on it
        uint32_t done = 0;
        while (!done)
            asm volatile(
                "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}\n"
                : "=r"(done)
                : "r"(bar_addr)
                : "memory");
    }
    uint32_t lt = 0xffffff00u + lane;  // idle lanes: a group of their own
    if (p < n_pods) {
        uint32_t pos[kMaxSlices];
#pragma unroll
        for (uint32_t s = 0; s < kMaxSlices; ++s) pos[s] = 0;
#pragma unroll
        for (uint32_t step = kSlice / 2; step > 0; step >>= 1) {
#pragma unroll
            for (uint32_t s = 0; s < kMaxSlices; ++s)
                if (s < n_slices && slices[s * kSlice + pos[s] + step - 1] < mine) pos[s] += step;
        }
        lt = 0;
#pragma unroll
        for (uint32_t s = 0; s < kMaxSlices; ++s)
            if (s < n_slices) lt += pos[s] + (slices[s * kSlice + pos[s]] < mine ? 1u : 0u);
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
// by binary search.  Every block first brings the sorted requests (<= 96 KB) into shared memory with one bulk
// asynchronous copy: fourteen dependent loads from shared memory instead of from L2.  The threads also re-arm the
// duplicate counters for the next call.
__global__ void __launch_bounds__(256)
k_node_rank(const u64* __restrict__ sorted, uint32_t n_pods, uint32_t* __restrict__ dup, const u64* __restrict__ used,
            const u64* __restrict__ cap, const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes, uint32_t dev_stride,
            RankDsts dst, const uint16_t* __restrict__ cut_src, uint16_t* __restrict__ cut_dst) {
    extern __shared__ __align__(16) u64 s_sorted[];
    __shared__ __align__(8) uint64_t bar;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    pdl_launch_dependents();
    const uint32_t bar_addr = (uint32_t)__cvta_generic_to_shared(&bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_addr) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    bool eligible = false;
    const u64 key = node_key(used, cap, nodeinfo, n_nodes, dev_stride, i, &eligible);  // independent of the request side
    pdl_wait();  // the sorted requests
    if (threadIdx.x == 0) {
        const uint32_t bytes = ((n_pods + 1u) & ~1u) * 8u;  // whole 16-byte units (the array has room for GS_MAX_BATCH)
        asm volatile("fence.proxy.async;" ::: "memory");    // the copy reads what the previous kernel's threads wrote
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_addr), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         (uint32_t)__cvta_generic_to_shared(s_sorted)),
                     "l"(sorted), "r"(bytes), "r"(bar_addr)
                     : "memory");
    }
    if (i < n_pods) dup[i] = 0;
    // results that go to host memory leave from this kernel only (one flush over PCIe at its end, not one per kernel)
    if (cut_dst && i < n_pods) cut_dst[i] = cut_src[i];
    __syncthreads();  // the barrier is initialised before anybody waits on it
    {
        uint32_t done = 0;
        while (!done)
            asm volatile(
                "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}\n"
                : "=r"(done)
                : "r"(bar_addr)
                : "memory");
    }
    if (i >= dst.count) return;
    uint32_t lo = 0;
    if (eligible) {  // (false beyond the nodes: the padding of a shard's slice is written as zero)
        uint32_t len = n_pods;
        while (len > 0) {
            const uint32_t half = len >> 1;
            if (s_sorted[lo + half] <= key) {
                lo += half + 1;
                len -= half + 1;
            } else {
                len = half;
            }
        }
    }
    // one destination, or this shard's slice of the gathered array on every GPU (peer stores over NVLink)
    for (uint32_t k = 0; k < dst.n; ++k) dst.p[k][dst.offset + i] = (uint16_t)lo;
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
template <class... Params, class... Args>
static cudaError_t launch_pdl(void (*kernel)(Params...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // the launch itself overlaps the predecessor's tail
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<Params>(args)...);
}

cudaError_t launch_compact(const u64* d_req, uint32_t n_pods, u64* d_slices, u64* d_sorted, uint32_t* d_dup,
                           const u64* d_used, const u64* d_cap, const uint32_t* d_nodeinfo, uint32_t n_nodes,
                           uint32_t dev_stride, const RankDsts& dsts, uint16_t* d_cutoff, int sm_count, cudaStream_t st,
                           uint32_t* n_launches, u64* d_req_copy, uint16_t* d_cut_tmp) {
    *n_launches = 0;
    if (n_pods == 0 || n_pods > GS_MAX_BATCH) return cudaErrorInvalidValue;
    (void)sm_count;
    const uint32_t n_slices = (n_pods + kSlice - 1) / kSlice;
    // the first kernel of the chain is an ordinary launch: it waits for everything before it in the stream
    static bool attrs_done[64] = {};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64 || !attrs_done[dev]) {
        const int max_smem = (int)(kMaxSlices * kSlice * sizeof(u64));  // 96 KB: all slices
        cudaError_t ea = cudaFuncSetAttribute(k_req_place, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (ea != cudaSuccess) return ea;
        ea = cudaFuncSetAttribute(k_node_rank, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (ea != cudaSuccess) return ea;
        if (dev >= 0 && dev < 64) attrs_done[dev] = true;
    }
    // d_req_copy / d_cut_tmp: given when the requests and the cutoffs live in (page-locked) host memory
    k_slice_sort<<<n_slices, kSlice, 0, st>>>(d_req, n_pods, d_slices, d_req_copy);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    ++*n_launches;
    // small blocks: the searches of a pod are latency, so they are spread over as many SMs as there are blocks
    e = launch_pdl(k_req_place, dim3((n_pods + 63) / 64), dim3(64), (size_t)n_slices * kSlice * sizeof(u64), st,
                   d_req_copy ? (const u64*)nullptr : d_req, (const u64*)d_req_copy, n_pods, n_slices, (const u64*)d_slices, d_dup,
                   d_sorted, d_cut_tmp ? d_cut_tmp : d_cutoff);
    if (e != cudaSuccess) return e;
    ++*n_launches;
    const uint32_t n_threads = dsts.count > n_pods ? dsts.count : n_pods;
    e = launch_pdl(k_node_rank, dim3((n_threads + 255) / 256), dim3(256), (size_t)((n_pods + 1u) & ~1u) * sizeof(u64), st, (const u64*)d_sorted,
                   n_pods, d_dup, d_used, d_cap, d_nodeinfo, n_nodes, dev_stride, dsts, (const uint16_t*)d_cut_tmp,
                   d_cut_tmp ? d_cutoff : (uint16_t*)nullptr);
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
