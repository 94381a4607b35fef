// K3 -- batched Filter: the Predicate.Handler loop (pkg/scheduler/predicate.go:15-39) over
// NodeInfo.Assume (pkg/cache/nodeinfo.go:113-137) for P pending pods x N nodes.
//
// The predicate for (pod p, node n) is   class[n]==OK  &&  ndev[n]>0  &&  max_d(cap[n]-used[n][d]) >= req[p]
// (u64, wrapping subtraction: nodeinfo.go:260).  It is monotone in req[p], so with the pods sorted by
// request every node's column is a step: ones up to the last request <= key[n], zeros after.  An exact
// sort is not needed: it is enough to group the pods into narrow value buckets and settle the nodes
// whose key falls inside a tile's own bucket range by direct comparison with each row's
// request.  Two kernels:
//   k_prep   : pod blocks group the (<= GS_MAX_BATCH) requests into <= 512 value buckets; node blocks compute,
//              per node, key = max available over its devices and the bucket the key falls in, bucket ids
//              written bit-sliced (10 planes per 32-node word).
//   k_expand : writes the [P][N/32] bitmap tile by tile.  One thread turns the bucket ids of 32 nodes into
//              a word with 10 LOP3s; only nodes whose bucket lies inside the tile's own bucket range need
//              the exact comparison.
// (k_radix_sort, a stable single-CTA LSD radix sort, orders the bind queue by node for K4.)
#include "common.cuh"
#include "kernels.h"

namespace gs {

using u64 = unsigned long long;

// ------------------------------------------------------------------------------------------------
// single-CTA stable LSD radix sort of n <= GS_MAX_BATCH 64-bit keys.  Only the bit span in which the
// keys differ is sorted, in ceil(span / 9) passes of equal width (two 9-bit passes for MiB-sized
// requests, eight 8-bit passes for arbitrary 64-bit keys).  Ranking inside a warp is the ballot
// multi-split (one ballot per digit bit), per-warp counters, no atomics.
// ------------------------------------------------------------------------------------------------
constexpr int kSortThreads = 1024;
constexpr int kSortWarps = kSortThreads / 32;
constexpr int kSortMaxRounds = GS_MAX_BATCH / kSortThreads;  // rounds of 32 keys per warp
constexpr int kSortMaxBits = 9;
constexpr int kSortBins = 1 << kSortMaxBits;
constexpr size_t kSortSmemBytes = (size_t)GS_MAX_BATCH * 8 + (size_t)GS_MAX_BATCH * 2 * 2 +
                                  (size_t)kSortBins * kSortWarps * 2 + kSortBins * 2 + 128 * 8;
static_assert(GS_MAX_BATCH % kSortThreads == 0, "batch limit must be a whole number of rounds");

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
    uint16_t* hist = idx_b + GS_MAX_BATCH;            // [warp][bin]
    uint16_t* bin_base = hist + kSortBins * kSortWarps;  // [bin]
    u64* red = reinterpret_cast<u64*>(bin_base + kSortBins);

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
    if (varying != 0) {
        const int lo = __ffsll((long long)varying) - 1, hi = 63 - __clzll((long long)varying);
        const int span = hi - lo + 1;
        const int passes = (span + kSortMaxBits - 1) / kSortMaxBits;
        const int bits = (span + passes - 1) / passes;
        const uint32_t rounds = (n + kSortThreads - 1) / kSortThreads;
        const uint32_t chunk = rounds * 32;  // consecutive positions per warp
        const uint32_t lt_mask = (1u << lane) - 1u;
        uint16_t* my_hist = hist + warp * kSortBins;
        uint32_t* wsum = reinterpret_cast<uint32_t*>(red + 65);

        for (int pass = 0; pass < passes; ++pass) {
            const int shift = lo + pass * bits;
            const int nb = min(bits, hi + 1 - shift);
            const uint32_t nbins = 1u << nb, dmask = nbins - 1u;

            for (uint32_t i = lane; i < kSortBins / 2; i += 32) reinterpret_cast<uint32_t*>(my_hist)[i] = 0;
            __syncwarp();

            // rank every key among the keys of its warp's chunk that share its digit
            uint32_t slot[kSortMaxRounds];  // (digit << 16) | offset within (warp, digit)
#pragma unroll
            for (int r = 0; r < kSortMaxRounds; ++r) {
                slot[r] = 0;
                if ((uint32_t)r < rounds) {
                    const uint32_t pos = warp * chunk + r * 32 + lane;
                    const bool valid = pos < n;
                    uint32_t digit = 0;
                    if (valid) digit = (uint32_t)(keys[src[pos]] >> shift) & dmask;
                    uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
                    for (int b = 0; b < kSortMaxBits; ++b) {
                        if (b < nb) {
                            const bool bit = (digit >> b) & 1u;
                            const uint32_t bal = __ballot_sync(0xffffffffu, bit);
                            peers &= bit ? bal : ~bal;
                        }
                    }
                    uint32_t base = 0;
                    int leader = 0;
                    if (valid) {
                        leader = __ffs(peers) - 1;
                        if ((int)lane == leader) {
                            base = my_hist[digit];
                            my_hist[digit] = (uint16_t)(base + __popc(peers));
                        }
                    }
                    base = __shfl_sync(0xffffffffu, base, leader);
                    slot[r] = (digit << 16) | (base + __popc(peers & lt_mask));
                    __syncwarp();
                }
            }
            __syncthreads();

            // per bin: exclusive prefix over warps (in place) and the bin total
            uint32_t total = 0;
            if (tid < nbins) {
#pragma unroll 8
                for (int w = 0; w < kSortWarps; ++w) {
                    const uint32_t c = hist[w * kSortBins + tid];
                    hist[w * kSortBins + tid] = (uint16_t)total;
                    total += c;
                }
            }
            // exclusive scan of the bin totals over the block
            uint32_t incl = total;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if ((int)lane >= o) incl += t;
            }
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
            if (tid < nbins) bin_base[tid] = (uint16_t)(wsum[warp] + incl - total);
            __syncthreads();

#pragma unroll
            for (int r = 0; r < kSortMaxRounds; ++r) {
                if ((uint32_t)r < rounds) {
                    const uint32_t pos = warp * chunk + r * 32 + lane;
                    if (pos < n) {
                        const uint32_t digit = slot[r] >> 16, off = slot[r] & 0xffffu;
                        dst[(uint32_t)bin_base[digit] + my_hist[digit] + off] = src[pos];
                    }
                }
            }
            __syncthreads();
            uint16_t* t = src;
            src = dst;
            dst = t;
        }
    }

    for (uint32_t i = tid; i < n; i += kSortThreads) {
        const uint16_t j = src[i];
        out_perm[i] = j;
        out_keys[i] = keys[j];
    }
}

// ------------------------------------------------------------------------------------------------
// k_prep: pod side and node side in one launch (independent roles chosen by blockIdx)
// ------------------------------------------------------------------------------------------------
// Values are mapped to kBuckets value buckets, bucket(x) = min(x >> shift, kBuckets - 1), with `shift`
// derived on the host from the largest device capacity ever upserted, so requests up to that capacity
// spread over all buckets; the last bucket is open-ended.  Bucket ids are stored 1-based.
//
// Pod blocks: group the pending pods by bucket: every pod takes an arrival number in its bucket with an
// L2 atomic; after a barrier among the pod blocks each block turns the bucket counts into offsets and
// writes the grouped position of its own pods.  Order inside a bucket is whatever the atomics give; the
// bitmap does not depend on it.
//
// Node blocks: one warp per 32-node word, lane = node.  key = max available memory over the node's
// devices (nodeinfo.go:254-264, u64 wrap-around included) and
//   q' = 0                                   class != OK or no devices (nodeinfo.go:123)
//   q' = 1 + min(key >> shift, kBuckets)     otherwise
// so that for a closed bucket b'  q' > b' implies key >= every request in it and q' < b' implies
// key < all of them.  q' is written bit-sliced, plane-major inside blocks of kTileWords words:
//   planes[((word / kTileWords) * kRankPlanes + k) * kTileWords + word % kTileWords] = bit k of 32 q'.
#ifdef GS_EXPAND_TRACE
// development builds only: per launch (ring of 64), per kernel (0 = k_prep, 1 = k_expand, 2 = k_settle), earliest / latest
// globaltimer of: 0 block entry, 1 past the start wait, 2 block exit (tools/overlap_trace.py)
__device__ unsigned long long g_kmin[64][3][6];
__device__ unsigned long long g_kmax[64][3][6];
__device__ __forceinline__ void ktrace(uint32_t seq, int kernel, int what) {
    unsigned long long gt;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt));
    atomicMin(&g_kmin[seq & 63][kernel][what], gt);
    atomicMax(&g_kmax[seq & 63][kernel][what], gt);
}
#define GS_KTRACE(seq, kernel, what) do { if (threadIdx.x == 0) ktrace(seq, kernel, what); } while (0)
#else
#define GS_KTRACE(seq, kernel, what) do {} while (0)
#endif

// Epoch counters let consecutive filter batches overlap without stream-order waits: a kernel publishes "done"
// by bumping a counter after a fence, a later kernel that really depends on it spins on the counter.
__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void wait_epoch(const uint32_t* p, uint32_t want) {
    while ((int32_t)(ld_acquire_u32(p) - want) < 0) {
    }
}

constexpr int kPrepThreads = 128;  // small blocks: they fit next to a fully resident k_expand of the previous batch
constexpr uint32_t kBuckets = 512;

__device__ __forceinline__ u64 node_key_generic(const u64* __restrict__ used_row, u64 cap, uint32_t ndev) {
    u64 key = 0;
    for (uint32_t d = 0; d < ndev; ++d) {
        u64 a = cap - used_row[d];
        key = a > key ? a : key;
    }
    return key;
}

// Both roles read their inputs first and hold their results in registers; `may_write()` is called (by every
// thread of the block) before the first store into the buffer set.
template <int kDevStride, class MayWrite>
__device__ __forceinline__ void prep_node_role(uint32_t block, const FilterBuffers& fb, const u64* __restrict__ used,
                                               const u64* __restrict__ cap, const uint32_t* __restrict__ nodeinfo,
                                               uint32_t n_nodes, uint32_t dev_stride, uint32_t n_words, uint32_t shift,
                                               MayWrite may_write) {
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t word = block * (kPrepThreads / 32) + warp;
    const uint32_t node = word * 32 + lane;
    uint32_t q = 0;
    u64 key = 0;
    if (word < n_words && node < n_nodes) {
        const uint32_t info = nodeinfo[node];
        const uint32_t ndev = nodeinfo_ndev(info);
        if (nodeinfo_class(info) == GS_NODE_OK && ndev > 0) {
            const u64 c = cap[node];
            if (kDevStride == 8) {
                const ulonglong2* row = reinterpret_cast<const ulonglong2*>(used + (size_t)node * 8);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    ulonglong2 u = row[k];
                    u64 a0 = c - u.x, a1 = c - u.y;
                    if ((uint32_t)(2 * k) < ndev) key = a0 > key ? a0 : key;
                    if ((uint32_t)(2 * k + 1) < ndev) key = a1 > key ? a1 : key;
                }
            } else {
                key = node_key_generic(used + (size_t)node * dev_stride, c, ndev);
            }
            const u64 bq = key >> shift;
            q = 1u + (uint32_t)(bq < kBuckets ? bq : kBuckets);
        }
    }
    uint32_t mine = 0;
#pragma unroll
    for (int k = 0; k < kRankPlanes; ++k) {
        uint32_t b = __ballot_sync(0xffffffffu, (q >> k) & 1u);
        if ((int)lane == k) mine = b;
    }
    may_write();
    if (word >= n_words) return;
    fb.node_key[node] = key;
    fb.node_key32[node] = key > 0xffffffffull ? 0xffffffffu : (uint32_t)key;  // exact against any request below 2^32
    if (lane < kRankPlanes)
        fb.planes[((size_t)(word / kTileWords) * kRankPlanes + lane) * kTileWords + (word % kTileWords)] = mine;
}

template <int kDevStride>
__global__ void __launch_bounds__(kPrepThreads)
k_prep(const u64* __restrict__ req, uint32_t n_pods, uint32_t pod_blocks, uint32_t shift, FilterBuffers fb,
       const u64* __restrict__ used, const u64* __restrict__ cap, const uint32_t* __restrict__ nodeinfo,
       uint32_t n_nodes, uint32_t dev_stride, uint32_t n_words, uint32_t overlap_previous,
       uint32_t expand_epoch_wanted) {
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    __shared__ bool is_last;
    // Programmatic dependent launch, both ways.  This grid is launched as a programmatic dependent of whatever
    // precedes it in the stream: normally it waits for that work to complete and flush.  Only when the host knows
    // that the predecessor is the expand kernel of another filter batch, which shares nothing with this one, does
    // it start at once; what it must not do then is overwrite work buffers that an older expand kernel (the one
    // that last used this buffer set) may still be reading.  That kernel's epoch is read now and checked just
    // before the first store into the set, so the round trip overlaps the loads below (a round trip to L2
    // costs microseconds on an SM that is streaming the previous batch's bitmap).
    // k_expand may start its prologue as soon as every block here has started; it waits for this grid's own
    // epoch before it reads anything written here.
    GS_KTRACE(2 * (expand_epoch_wanted + 1) + fb.set_index, 0, 0);
    uint32_t epoch_seen = expand_epoch_wanted;
    if (!overlap_previous) {
        asm volatile("griddepcontrol.wait;" ::: "memory");
    } else if (tid == 0) {
        epoch_seen = ld_acquire_u32(fb.expand_epoch);
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    auto may_write = [&]() {
        if (tid == 0 && (int32_t)(epoch_seen - expand_epoch_wanted) < 0) wait_epoch(fb.expand_epoch, expand_epoch_wanted);
        __syncthreads();
        GS_KTRACE(2 * (expand_epoch_wanted + 1) + fb.set_index, 0, 1);
    };
    if (blockIdx.x < pod_blocks) {
        // ---------------- pod role ----------------
        __shared__ uint32_t base[kBuckets];
        __shared__ uint32_t wsum[kPrepThreads / 32];
        constexpr int kPer = kBuckets / kPrepThreads;  // buckets scanned per thread
        static_assert(kPer == 4, "the scan below reads four counts per thread");
        uint32_t* ghist = fb.pod_hist;  // [kBuckets] bucket counts, zero between launches
        const uint32_t i = blockIdx.x * kPrepThreads + tid;
        uint32_t sl = 0;
        u64 r = 0;
        if (i < n_pods) {
            r = req[i];
            const u64 bq = r >> shift;
            const uint32_t b = bq < kBuckets - 1 ? (uint32_t)bq : kBuckets - 1;
            sl = (b << 16) | atomicAdd(&ghist[b], 1u);  // arrival number inside the bucket
        }
        if (sl == 0xffffffffu) return;  // (never true: makes the stamp below wait for the atomic's result)
        GS_KTRACE(2 * (expand_epoch_wanted + 1) + fb.set_index, 0, 3);
        // The counts are complete when they add up to the number of pods: every block polls the 512 counts (the
        // pod blocks are the first blocks of a small grid of small blocks: all of them are resident), and the
        // poll that finds them complete already holds the values the scan needs.
        uint4 c4;
        uint32_t run;
        for (;;) {
            c4 = __ldcg(reinterpret_cast<const uint4*>(ghist) + tid);
            const uint32_t sum = c4.x + c4.y + c4.z + c4.w;
            uint32_t incl = sum;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                if ((int)lane >= o) incl += t;
            }
            if (lane == 31) wsum[warp] = incl;
            __syncthreads();
            run = incl - sum;
            uint32_t total = 0;
#pragma unroll
            for (uint32_t w = 0; w < kPrepThreads / 32; ++w) {
                const uint32_t ws = wsum[w];
                total += ws;
                if (w < warp) run += ws;
            }
            __syncthreads();  // wsum is rewritten by the next poll
            if (total == n_pods) break;
        }
        GS_KTRACE(2 * (expand_epoch_wanted + 1) + fb.set_index, 0, 4);
        base[tid * 4 + 0] = run;
        base[tid * 4 + 1] = run + c4.x;
        base[tid * 4 + 2] = run + c4.x + c4.y;
        base[tid * 4 + 3] = run + c4.x + c4.y + c4.z;
        may_write();  // also orders the writes of `base` before the reads below
        if (i < n_pods) {
            const uint32_t b = sl >> 16;
            const uint32_t j = base[b] + (sl & 0xffffu);
            fb.perm[j] = (uint16_t)i;
            fb.grouped_req[j] = r;
            fb.row_bucket[j] = (uint16_t)(b + 1);  // bucket ids are stored 1-based
        }
    } else {
        prep_node_role<kDevStride>(blockIdx.x - pod_blocks, fb, used, cap, nodeinfo, n_nodes, dev_stride, n_words, shift,
                                   may_write);
    }
    // Done: the last block out clears the bucket counts for the next launch (every pod block has read them by
    // then) and publishes this launch; everything written above is fenced first.
    __syncthreads();
    GS_KTRACE(2 * (expand_epoch_wanted + 1) + fb.set_index, 0, 2);
    if (tid == 0) {
        __threadfence();
        is_last = atomicAdd(fb.prep_done, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (is_last) {
        reinterpret_cast<uint4*>(fb.pod_hist)[tid] = make_uint4(0, 0, 0, 0);
        __syncthreads();
        if (tid == 0) {
            *fb.prep_done = 0;
            __threadfence();
            atomicAdd(fb.prep_epoch, 1u);
            // A grid that did not wait at its start must still not complete before its predecessor does: later
            // work in the stream relies on completion being transitive.  One lingering thread is enough to hold
            // the grid open; every other block frees its SM resources at once.
            asm volatile("griddepcontrol.wait;" ::: "memory");
        }
    }
}

// ------------------------------------------------------------------------------------------------
// bitmap expansion
// ------------------------------------------------------------------------------------------------
// bit b of the result = (q' of node b >= t), q' given as bit planes a[0..kRankPlanes) (LSB first)
__device__ __forceinline__ uint32_t planes_ge(const uint32_t (&a)[kRankPlanes], uint32_t t) {
    uint32_t ge = 0xffffffffu;
#pragma unroll
    for (int k = 0; k < kRankPlanes; ++k) ge = ((t >> k) & 1u) ? (a[k] & ge) : (a[k] | ge);  // t is warp-uniform
    return ge;
}

#include "tile_kernels.inc"

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
static bool g_attr_done[64] = {};

static cudaError_t ensure_attrs(int device) {
    if (device >= 0 && device < 64 && g_attr_done[device]) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(k_radix_sort<BindKeyLoader>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kSortSmemBytes);
    if (e != cudaSuccess) return e;
    if (device >= 0 && device < 64) g_attr_done[device] = true;
    return cudaSuccess;
}

cudaError_t launch_sort_binds(int device, const gs_bind_req* d_binds, uint32_t n, u64* d_sorted,
                              uint16_t* d_perm, cudaStream_t st) {
    cudaError_t e = ensure_attrs(device);
    if (e != cudaSuccess) return e;
    BindKeyLoader ld{d_binds};
    k_radix_sort<BindKeyLoader><<<1, kSortThreads, kSortSmemBytes, st>>>(ld, n, d_sorted, d_perm);
    return cudaGetLastError();
}

cudaError_t launch_prep(const u64* d_req, uint32_t n_pods, bool do_pods, bool do_nodes, uint32_t shift,
                        const FilterBuffers& fb, const u64* d_used, const u64* d_cap, const uint32_t* d_nodeinfo,
                        uint32_t n_nodes, uint32_t dev_stride, uint32_t n_words, bool overlap_previous,
                        uint32_t expand_epoch_wanted, cudaStream_t st) {
    const uint32_t pod_blocks = do_pods ? (n_pods + kPrepThreads - 1) / kPrepThreads : 0u;
    const uint32_t node_blocks = do_nodes ? (n_words + kPrepThreads / 32 - 1) / (kPrepThreads / 32) : 0u;
    if (pod_blocks + node_blocks == 0) return cudaErrorInvalidValue;  // callers count launches: never a silent no-op
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(pod_blocks + node_blocks);
    cfg.blockDim = dim3(kPrepThreads);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const uint32_t ov = overlap_previous ? 1u : 0u;
    if (dev_stride == 8)
        return cudaLaunchKernelEx(&cfg, k_prep<8>, d_req, n_pods, pod_blocks, shift, fb, d_used, d_cap, d_nodeinfo, n_nodes,
                                  dev_stride, n_words, ov, expand_epoch_wanted);
    return cudaLaunchKernelEx(&cfg, k_prep<0>, d_req, n_pods, pod_blocks, shift, fb, d_used, d_cap, d_nodeinfo, n_nodes,
                              dev_stride, n_words, ov, expand_epoch_wanted);
}

// Shape of the tile grid for one launch; k_settle and k_expand must agree on it.
struct TileGeometry {
    bool vec;
    uint32_t words_out, n_col_blocks, n_row_tiles, n_tiles, inv_col_blocks;
};
static TileGeometry tile_geometry(uint32_t n_pods, uint32_t words, uint32_t row_stride, const uint32_t* d_out) {
    TileGeometry g;
    g.vec = (row_stride % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_out) & 15u) == 0);
    // 16-byte stores; the row's last 32-byte sector is written whole when the stride leaves room for it (a
    // half-written sector costs a read-modify-write in DRAM: profiles/NOTES.md).  Padding bits are zero.
    g.words_out = words;
    if (g.vec) {
        g.words_out = (words + 7u) & ~7u;
        if (g.words_out > row_stride) g.words_out = (words + 3u) & ~3u;
    }
    g.n_col_blocks = (g.words_out + kTileWords - 1) / kTileWords;
    g.n_row_tiles = (n_pods + kTileRows - 1) / kTileRows;
    g.n_tiles = g.n_col_blocks * g.n_row_tiles;
    // exact t / n_col_blocks as a multiply-high for every t below 2^32 / n_col_blocks (n_tiles is far
    // smaller); 0 stands for "one column block" (2^32 does not fit)
    g.inv_col_blocks = g.n_col_blocks == 1 ? 0u : (uint32_t)((0x100000000ull + g.n_col_blocks - 1) / g.n_col_blocks);
    return g;
}

size_t tile_record_bytes() { return sizeof(TileRecord); }
size_t tile_payload_bytes() { return sizeof(TilePayload); }

static void pdl_config(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attr, uint32_t grid, uint32_t block, cudaStream_t st) {
    cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // may start under the predecessor's tail
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
}

cudaError_t launch_settle(int sm_count, const FilterBuffers& fb, uint32_t n_pods, uint32_t words, uint32_t row_stride,
                          const uint32_t* d_out, bool overlap_previous, uint32_t prep_epoch_wanted, cudaStream_t st) {
    if (n_pods == 0 || words == 0) return cudaErrorInvalidValue;  // callers count launches: never a silent no-op
    const TileGeometry g = tile_geometry(n_pods, words, row_stride, d_out);
    // about four CTAs per SM; each takes every n_splits-th row tile of one column block
    const uint32_t want = (uint32_t)sm_count * 4u;
    uint32_t splits = want / g.n_col_blocks;
    const uint32_t min_splits = (g.n_row_tiles + kSettleRowTiles - 1) / kSettleRowTiles;
    if (splits < min_splits) splits = min_splits;
    if (splits > g.n_row_tiles) splits = g.n_row_tiles;
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    pdl_config(cfg, attr, g.n_col_blocks * splits, kSettleThreads, st);
    return cudaLaunchKernelEx(&cfg, k_settle, fb, n_pods, g.n_col_blocks, g.n_row_tiles, splits,
                              overlap_previous ? 1u : 0u, prep_epoch_wanted);
}

cudaError_t launch_expand(int sm_count, const FilterBuffers& fb, uint32_t n_pods, uint32_t words,
                          uint32_t row_stride, uint32_t* d_out, bool overlap_previous, uint32_t settle_epoch_wanted,
                          cudaStream_t st) {
    if (n_pods == 0 || words == 0) return cudaErrorInvalidValue;  // callers count launches: never a silent no-op
    const TileGeometry g = tile_geometry(n_pods, words, row_stride, d_out);
    // Four CTAs per SM stream as fast as five (tools/chunk_write_bw.cu) and leave registers and threads for
    // the small kernels of the next batch, all of whose blocks must be resident next to a running k_expand
    // (a programmatic dependent is launched only when ALL blocks of its primary have started).
    // The limit holds across overlapping launches because it is set through shared memory: static (6 KB) + this
    // unused dynamic pad + 1 KB reserved > 228 KB / 5, which still leaves ~40 KB per SM for k_prep and k_settle.
    const uint32_t per_sm = 4;
    constexpr uint32_t kOccupancyPadBytes = 40 * 1024;
    uint32_t grid = (uint32_t)sm_count * per_sm;
    if (grid > g.n_tiles) grid = g.n_tiles;
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    pdl_config(cfg, attr, grid, kExpandThreads, st);
    cfg.dynamicSmemBytes = kOccupancyPadBytes;
    const uint32_t ov = overlap_previous ? 1u : 0u;
    if (g.vec)
        return cudaLaunchKernelEx(&cfg, k_expand<true>, fb, n_pods, g.words_out, row_stride, d_out, g.n_col_blocks,
                                  g.inv_col_blocks, g.n_tiles, ov, settle_epoch_wanted);
    return cudaLaunchKernelEx(&cfg, k_expand<false>, fb, n_pods, g.words_out, row_stride, d_out, g.n_col_blocks,
                              g.inv_col_blocks, g.n_tiles, ov, settle_epoch_wanted);
}

}  // namespace gs

#ifdef GS_EXPAND_TRACE
extern "C" int gs_debug_ktrace_reset() {
    static unsigned long long lo[64 * 3 * 6], hi[64 * 3 * 6];
    for (auto& v : lo) v = ~0ull;
    for (auto& v : hi) v = 0;
    cudaError_t e = cudaMemcpyToSymbol(gs::g_kmin, lo, sizeof lo);
    if (e == cudaSuccess) e = cudaMemcpyToSymbol(gs::g_kmax, hi, sizeof hi);
    return (int)e;
}
extern "C" int gs_debug_ktrace_read(unsigned long long* lo, unsigned long long* hi) {
    cudaError_t e = cudaMemcpyFromSymbol(lo, gs::g_kmin, sizeof(gs::g_kmin));
    if (e == cudaSuccess) e = cudaMemcpyFromSymbol(hi, gs::g_kmax, sizeof(gs::g_kmax));
    return (int)e;
}
#endif
