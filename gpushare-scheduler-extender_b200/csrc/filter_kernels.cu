// K3 -- batched Filter: the Predicate.Handler loop (pkg/scheduler/predicate.go:15-39) over
// NodeInfo.Assume (pkg/cache/nodeinfo.go:113-137) for P pending pods x N nodes.
//
// The predicate for (pod p, node n) is   class[n]==OK  &&  ndev[n]>0  &&  max_d(cap[n]-used[n][d]) >= req[p]
// (u64, wrapping subtraction: nodeinfo.go:260).  It is monotone in req[p], so with the pods sorted by
// request every node's column is a step: ones for the first rank[n] sorted rows, zeros after, where
// rank[n] = #{p : req[p] <= key[n]}.  Three kernels:
//   k_radix_sort : one CTA sorts the (<=16384) requests, LSD radix on the bits that actually vary.
//   k_node_rank  : per node, key = max available over its devices, rank = upper_bound(sorted req, key);
//                  ranks are written bit-sliced (16 planes per 32-node word).
//   k_expand     : writes the [P][N/32] bitmap, tile by tile, from the bit-sliced ranks.  One thread
//                  turns 32 ranks into a word with 16 LOP3s; rows of a tile differ from the tile's base
//                  word only in the few words that hold a step, which are patched in shared memory.
#include "common.cuh"
#include "kernels.h"

namespace gs {

using u64 = unsigned long long;

// ------------------------------------------------------------------------------------------------
// single-CTA stable LSD radix sort of n <= GS_MAX_BATCH 64-bit keys (8-bit digits)
// ------------------------------------------------------------------------------------------------
constexpr int kSortThreads = 1024;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortMaxRounds = GS_MAX_BATCH / kSortThreads;  // 16 rounds of 32 keys per warp
constexpr size_t kSortSmemBytes = (size_t)GS_MAX_BATCH * 8 + (size_t)GS_MAX_BATCH * 2 * 2 +
                                  256 * kSortWarps * 2 + 128 * 8;

struct ReqKeyLoader {
    const u64* req;
    __device__ u64 operator()(uint32_t i) const { return req[i]; }
};
struct BindKeyLoader {
    const gs_bind_req* binds;
    __device__ u64 operator()(uint32_t i) const { return (u64)binds[i].node_idx; }
};

template <class Loader>
__global__ void __launch_bounds__(kSortThreads, 1)
k_radix_sort(Loader load_key, uint32_t n, u64* __restrict__ out_keys, uint16_t* __restrict__ out_perm) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    u64* keys = reinterpret_cast<u64*>(smem_raw);
    uint16_t* idx_a = reinterpret_cast<uint16_t*>(keys + GS_MAX_BATCH);
    uint16_t* idx_b = idx_a + GS_MAX_BATCH;
    uint16_t* hist = idx_b + GS_MAX_BATCH;  // [256 digits][32 warps]
    u64* red = reinterpret_cast<u64*>(hist + 256 * kSortWarps);

    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    u64 v_or = 0, v_and = ~0ull;
    for (uint32_t i = tid; i < n; i += kSortThreads) {
        u64 k = load_key(i);
        keys[i] = k;
        idx_a[i] = (uint16_t)i;
        v_or |= k;
        v_and &= k;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v_or |= __shfl_xor_sync(0xffffffffu, v_or, o);
        v_and &= __shfl_xor_sync(0xffffffffu, v_and, o);
    }
    if (lane == 0) {
        red[warp] = v_or;
        red[32 + warp] = v_and;
    }
    __syncthreads();
    if (warp == 0) {
        u64 a = red[lane], b = red[32 + lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a |= __shfl_xor_sync(0xffffffffu, a, o);
            b &= __shfl_xor_sync(0xffffffffu, b, o);
        }
        if (lane == 0) red[64] = a ^ b;
    }
    __syncthreads();
    const u64 varying = red[64];

    uint16_t* src = idx_a;
    uint16_t* dst = idx_b;
    const uint32_t rounds = (n + kSortThreads - 1) / kSortThreads;  // rounds of 32 keys per warp
    const uint32_t chunk = rounds * 32;                             // consecutive positions per warp
    const uint32_t lt_mask = (1u << lane) - 1u;

    for (int shift = 0; shift < 64; shift += 8) {
        if (((varying >> shift) & 0xffull) == 0) continue;  // block-uniform: digit identical everywhere
        for (uint32_t i = tid; i < 256 * kSortWarps; i += kSortThreads) hist[i] = 0;
        __syncthreads();

        // rank every key among the keys of its warp's chunk that share its digit
        uint32_t slot[kSortMaxRounds];  // (digit << 16) | offset within (digit, warp)
#pragma unroll
        for (int r = 0; r < kSortMaxRounds; ++r) {
            slot[r] = 0;
            if ((uint32_t)r < rounds) {
                const uint32_t pos = warp * chunk + r * 32 + lane;
                const bool valid = pos < n;
                uint32_t digit = 0x1ffu;
                if (valid) digit = (uint32_t)(keys[src[pos]] >> shift) & 0xffu;
                const uint32_t peers = __match_any_sync(0xffffffffu, digit);
                const int leader = __ffs(peers) - 1;
                uint32_t base = 0;
                if ((int)lane == leader && valid) {
                    base = hist[digit * kSortWarps + warp];
                    hist[digit * kSortWarps + warp] = (uint16_t)(base + __popc(peers));
                }
                base = __shfl_sync(0xffffffffu, base, leader);
                slot[r] = (digit << 16) | (base + __popc(peers & lt_mask));
                __syncwarp();
            }
        }
        __syncthreads();

        // exclusive scan of hist in (digit, warp) order: 8 consecutive counters per thread
        {
            uint4 raw = *reinterpret_cast<uint4*>(hist + tid * 8);
            uint32_t c[8] = {raw.x & 0xffffu, raw.x >> 16, raw.y & 0xffffu, raw.y >> 16,
                             raw.z & 0xffffu, raw.z >> 16, raw.w & 0xffffu, raw.w >> 16};
            uint32_t sum = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) sum += c[i];
            uint32_t incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if ((int)lane >= o) incl += t;
            }
            uint32_t* wsum = reinterpret_cast<uint32_t*>(red + 65);
            if (lane == 31) wsum[warp] = incl;
            __syncthreads();
            if (warp == 0) {
                uint32_t w = wsum[lane], wi = w;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
                    if ((int)lane >= o) wi += t;
                }
                wsum[lane] = wi - w;
            }
            __syncthreads();
            uint32_t run = wsum[warp] + incl - sum;
            uint32_t e[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                e[i] = run;
                run += c[i];
            }
            raw.x = e[0] | (e[1] << 16);
            raw.y = e[2] | (e[3] << 16);
            raw.z = e[4] | (e[5] << 16);
            raw.w = e[6] | (e[7] << 16);
            *reinterpret_cast<uint4*>(hist + tid * 8) = raw;
        }
        __syncthreads();

#pragma unroll
        for (int r = 0; r < kSortMaxRounds; ++r) {
            if ((uint32_t)r < rounds) {
                const uint32_t pos = warp * chunk + r * 32 + lane;
                if (pos < n) {
                    const uint32_t digit = slot[r] >> 16, off = slot[r] & 0xffffu;
                    dst[hist[digit * kSortWarps + warp] + off] = src[pos];
                }
            }
        }
        __syncthreads();
        uint16_t* t = src;
        src = dst;
        dst = t;
    }

    for (uint32_t i = tid; i < n; i += kSortThreads) {
        const uint16_t j = src[i];
        out_perm[i] = j;
        out_keys[i] = keys[j];
    }
}

// ------------------------------------------------------------------------------------------------
// per-node key and rank, written as bit planes
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ u64 node_key_generic(const u64* __restrict__ used_row, u64 cap, uint32_t ndev) {
    u64 key = 0;
    for (uint32_t d = 0; d < ndev; ++d) {
        u64 a = cap - used_row[d];
        key = a > key ? a : key;
    }
    return key;
}

// One warp per 32-node word; lane = node.
template <int kDevStride>
__global__ void __launch_bounds__(256)
k_node_rank(const u64* __restrict__ used, const u64* __restrict__ cap, const uint32_t* __restrict__ nodeinfo,
            uint32_t n_nodes, uint32_t dev_stride, const u64* __restrict__ sorted_req, uint32_t n_pods,
            uint32_t* __restrict__ planes, uint32_t n_words) {
    const uint32_t lane = threadIdx.x & 31;
    const uint32_t word = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (word >= n_words) return;
    const uint32_t node = word * 32 + lane;
    uint32_t rank = 0;
    if (node < n_nodes) {
        const uint32_t info = nodeinfo[node];
        const uint32_t ndev = nodeinfo_ndev(info);
        if (nodeinfo_class(info) == GS_NODE_OK && ndev > 0) {
            const u64 c = cap[node];
            u64 key = 0;
            if (kDevStride == 8) {
                const ulonglong2* row = reinterpret_cast<const ulonglong2*>(used + (size_t)node * 8);
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    ulonglong2 u = row[q];
                    u64 a0 = c - u.x, a1 = c - u.y;
                    if ((uint32_t)(2 * q) < ndev) key = a0 > key ? a0 : key;
                    if ((uint32_t)(2 * q + 1) < ndev) key = a1 > key ? a1 : key;
                }
            } else {
                key = node_key_generic(used + (size_t)node * dev_stride, c, ndev);
            }
            // upper_bound: number of sorted requests <= key
            uint32_t lo = 0, hi = n_pods;
            while (lo < hi) {
                uint32_t mid = (lo + hi) >> 1;
                if (__ldg(sorted_req + mid) <= key)
                    lo = mid + 1;
                else
                    hi = mid;
            }
            rank = lo;
        }
    }
    uint32_t mine = 0;
#pragma unroll
    for (int k = 0; k < kRankPlanes; ++k) {
        uint32_t b = __ballot_sync(0xffffffffu, (rank >> k) & 1u);
        if ((int)lane == k) mine = b;
    }
    if (lane < kRankPlanes) planes[(size_t)word * kRankPlanes + lane] = mine;
}

// ------------------------------------------------------------------------------------------------
// bitmap expansion
// ------------------------------------------------------------------------------------------------
// bit b of the result = (rank of node b >= t), ranks given as bit planes a[0..15] (LSB first)
__device__ __forceinline__ uint32_t planes_ge(const uint32_t (&a)[kRankPlanes], uint32_t t) {
    uint32_t ge = 0xffffffffu;
#pragma unroll
    for (int k = 0; k < kRankPlanes; ++k) {
        const uint32_t m = 0u - ((t >> k) & 1u);
        ge = (m & a[k] & ge) | (~m & (a[k] | ge));
    }
    return ge;
}

struct ExpandSmem {
    uint32_t tile[kTileRows * kTileStride];
    uint32_t planes[kTileWords * kRankPlanes];
    uint32_t basew[kTileWords];
    uint32_t row_of[kTileRows];
    uint32_t plist[kTileWords];
    uint32_t n_partial;
};

template <bool kVec>
__global__ void __launch_bounds__(kExpandThreads)
k_expand(const uint32_t* __restrict__ planes, const uint16_t* __restrict__ perm, uint32_t n_pods,
         uint32_t words_out, uint32_t row_stride, uint32_t* __restrict__ out, uint32_t n_col_blocks,
         uint32_t n_tiles) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    ExpandSmem& sm = *reinterpret_cast<ExpandSmem*>(smem_raw);
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr uint32_t kWarps = kExpandThreads / 32;

    for (uint32_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint32_t rt = tile / n_col_blocks, cb = tile - rt * n_col_blocks;
        const uint32_t j0 = rt * kTileRows;
        const uint32_t rows = min((uint32_t)kTileRows, n_pods - j0);
        const uint32_t j1 = j0 + rows;
        const uint32_t w0 = cb * kTileWords;

        if (tid == 0) sm.n_partial = 0;
        __syncthreads();

        // A: one thread per word: base word (nodes whose step lies at or beyond the tile's last row)
        //    and the list of words that hold a step inside the tile
        if (tid < kTileWords) {
            uint32_t a[kRankPlanes];
            const uint4* src = reinterpret_cast<const uint4*>(planes + (size_t)(w0 + tid) * kRankPlanes);
#pragma unroll
            for (int q = 0; q < kRankPlanes / 4; ++q) {
                uint4 v = __ldg(src + q);
                a[4 * q] = v.x;
                a[4 * q + 1] = v.y;
                a[4 * q + 2] = v.z;
                a[4 * q + 3] = v.w;
                *reinterpret_cast<uint4*>(&sm.planes[tid * kRankPlanes + 4 * q]) = v;
            }
            const uint32_t all_rows = planes_ge(a, j1);     // rank >= j1: set in every row of the tile
            const uint32_t some_rows = planes_ge(a, j0 + 1);  // rank >  j0: set in the first row at least
            sm.basew[tid] = all_rows;
            if (some_rows & ~all_rows) sm.plist[atomicAdd(&sm.n_partial, 1u)] = tid;
        } else if (tid < kTileWords + kTileRows) {
            const uint32_t r = tid - kTileWords;
            sm.row_of[r] = r < rows ? (uint32_t)perm[j0 + r] : 0u;
        }
        __syncthreads();

        // B: fill the tile with the base words
        {
            const uint4 b4 = *reinterpret_cast<const uint4*>(&sm.basew[lane * 4]);
            for (uint32_t r = warp; r < rows; r += kWarps)
                *reinterpret_cast<uint4*>(&sm.tile[r * kTileStride + lane * 4]) = b4;
        }
        __syncthreads();

        // C: words with a step inside the tile: lane = row, exact word from the planes
        {
            const uint32_t np = sm.n_partial;
            for (uint32_t k = warp; k < np; k += kWarps) {
                const uint32_t t = sm.plist[k];
                uint32_t a[kRankPlanes];
#pragma unroll
                for (int q = 0; q < kRankPlanes; ++q) a[q] = sm.planes[t * kRankPlanes + q];
                for (uint32_t r = lane; r < rows; r += 32)
                    sm.tile[r * kTileStride + t] = planes_ge(a, j0 + r + 1);
            }
        }
        __syncthreads();

        // D: rows go to their pods' rows of the output
        if (kVec) {
            const uint32_t wcol = w0 + lane * 4;
            if (wcol < words_out) {
                for (uint32_t r = warp; r < rows; r += kWarps) {
                    const uint4 v = *reinterpret_cast<const uint4*>(&sm.tile[r * kTileStride + lane * 4]);
                    __stcs(reinterpret_cast<uint4*>(out + (size_t)sm.row_of[r] * row_stride + wcol), v);
                }
            }
        } else {
            for (uint32_t r = warp; r < rows; r += kWarps) {
                uint32_t* dst = out + (size_t)sm.row_of[r] * row_stride;
#pragma unroll
                for (int c = 0; c < kTileWords / 32; ++c) {
                    const uint32_t wl = c * 32 + lane;
                    if (w0 + wl < words_out) __stcs(dst + w0 + wl, sm.tile[r * kTileStride + wl]);
                }
            }
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
static bool g_attr_done[64] = {};

static cudaError_t ensure_attrs(int device) {
    if (device >= 0 && device < 64 && g_attr_done[device]) return cudaSuccess;
    cudaError_t e;
    e = cudaFuncSetAttribute(k_radix_sort<ReqKeyLoader>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)kSortSmemBytes);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_radix_sort<BindKeyLoader>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)kSortSmemBytes);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_expand<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ExpandSmem));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_expand<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ExpandSmem));
    if (e != cudaSuccess) return e;
    if (device >= 0 && device < 64) g_attr_done[device] = true;
    return cudaSuccess;
}

cudaError_t launch_sort_requests(int device, const u64* d_req, uint32_t n, u64* d_sorted, uint16_t* d_perm,
                                 cudaStream_t st) {
    cudaError_t e = ensure_attrs(device);
    if (e != cudaSuccess) return e;
    ReqKeyLoader ld{d_req};
    k_radix_sort<ReqKeyLoader><<<1, kSortThreads, kSortSmemBytes, st>>>(ld, n, d_sorted, d_perm);
    return cudaGetLastError();
}

cudaError_t launch_sort_binds(int device, const gs_bind_req* d_binds, uint32_t n, u64* d_sorted,
                              uint16_t* d_perm, cudaStream_t st) {
    cudaError_t e = ensure_attrs(device);
    if (e != cudaSuccess) return e;
    BindKeyLoader ld{d_binds};
    k_radix_sort<BindKeyLoader><<<1, kSortThreads, kSortSmemBytes, st>>>(ld, n, d_sorted, d_perm);
    return cudaGetLastError();
}

cudaError_t launch_node_rank(const u64* d_used, const u64* d_cap, const uint32_t* d_nodeinfo, uint32_t n_nodes,
                             uint32_t dev_stride, const u64* d_sorted, uint32_t n_pods, uint32_t* d_planes,
                             uint32_t n_words, cudaStream_t st) {
    const uint32_t threads = 256;
    const uint32_t blocks = (n_words * 32 + threads - 1) / threads;
    if (blocks == 0) return cudaSuccess;
    if (dev_stride == 8)
        k_node_rank<8><<<blocks, threads, 0, st>>>(d_used, d_cap, d_nodeinfo, n_nodes, dev_stride, d_sorted,
                                                    n_pods, d_planes, n_words);
    else
        k_node_rank<0><<<blocks, threads, 0, st>>>(d_used, d_cap, d_nodeinfo, n_nodes, dev_stride, d_sorted,
                                                    n_pods, d_planes, n_words);
    return cudaGetLastError();
}

cudaError_t launch_expand(int device, int sm_count, const uint32_t* d_planes, const uint16_t* d_perm,
                          uint32_t n_pods, uint32_t words, uint32_t row_stride, uint32_t* d_out,
                          cudaStream_t st) {
    cudaError_t e = ensure_attrs(device);
    if (e != cudaSuccess) return e;
    if (n_pods == 0 || words == 0) return cudaSuccess;
    const bool vec = (row_stride % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_out) & 15u) == 0);
    const uint32_t words_out = vec ? ((words + 3u) & ~3u) : words;
    const uint32_t n_col_blocks = (words_out + kTileWords - 1) / kTileWords;
    const uint32_t n_row_tiles = (n_pods + kTileRows - 1) / kTileRows;
    const uint32_t n_tiles = n_col_blocks * n_row_tiles;
    const uint32_t per_sm = 4;
    uint32_t grid = (uint32_t)sm_count * per_sm;
    if (grid > n_tiles) grid = n_tiles;
    if (vec)
        k_expand<true><<<grid, kExpandThreads, sizeof(ExpandSmem), st>>>(d_planes, d_perm, n_pods, words_out,
                                                                         row_stride, d_out, n_col_blocks, n_tiles);
    else
        k_expand<false><<<grid, kExpandThreads, sizeof(ExpandSmem), st>>>(d_planes, d_perm, n_pods, words_out,
                                                                          row_stride, d_out, n_col_blocks, n_tiles);
    return cudaGetLastError();
}

}  // namespace gs
