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
// development builds only: per launch (ring of 64), per kernel (0 = k_prep, 1 = k_expand), earliest / latest
// globaltimer of: 0 block entry, 1 past the start wait, 2 block exit (tools/overlap_trace.py)
__device__ unsigned long long g_kmin[64][2][6];
__device__ unsigned long long g_kmax[64][2][6];
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

// Every wait on another kernel's progress is bounded.  The waits below rest on assumptions about the launches
// around them (the awaited grid is resident or will become resident; counters are zero between launches); if one
// of them is ever broken -- a launch that was aborted and left its counters behind, a kernel of this library
// driven from outside the entry points -- the waiter says where it was stuck (a word in mapped host memory that
// the host reads when the next call fails) and traps, which ends the kernel with an error instead of holding
// the GPU for ever.
constexpr unsigned long long kSpinLimitNs = 4000000000ull;  // far above anything a healthy wait can last
enum : unsigned int {  // what gs_last_error() reports after such a trap
    kFaultPrepEpoch = 1,    // k_prep waiting for the expand kernel that last read its buffer set
    kFaultPodCounts = 2,    // k_prep pod blocks waiting for the other pod blocks' bucket counts
    kFaultExpandEpoch = 3,  // k_expand waiting for its own batch's k_prep
    kFaultTileInputs = 4,   // k_expand waiting for a bulk copy of tile inputs
    kFaultTileKeys = 5      // k_expand waiting for a bulk copy of node keys
};
__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __noinline__ void spin_give_up(unsigned int* fault, unsigned int where) {
    if (fault) {
        *reinterpret_cast<volatile unsigned int*>(fault) = where;
        __threadfence_system();
    }
    __trap();
}
struct SpinGuard {
    unsigned long long t0 = 0;
    uint32_t n = 0;
    __device__ __forceinline__ void tick(unsigned int* fault, unsigned int where) {
        if ((++n & 255u) != 0) return;
        const unsigned long long now = globaltimer_ns();
        if (t0 == 0)
            t0 = now;
        else if (now - t0 > kSpinLimitNs)
            spin_give_up(fault, where);
    }
};
__device__ __forceinline__ void wait_epoch(const uint32_t* p, uint32_t want, unsigned int* fault, unsigned int where) {
    SpinGuard g;
    while ((int32_t)(ld_acquire_u32(p) - want) < 0) {
#ifndef GS_NO_BACKOFF
        __nanosleep(256);  // thousands of waiting warps polling L2 flat out would slow the kernel they wait for
#endif
        g.tick(fault, where);
    }
}

// Overlapping batches: k_prep lets its dependent (the batch's k_expand) in only when a block has fenced its results,
// so a CTA of k_expand that starts knows them complete without reading an epoch first (0: k_prep triggers at its start
// and every CTA of k_expand polls the epoch before it asks for its first tile: one more round trip per CTA, 14.6 us per
// step instead of 13.9).
// (1: from a CTA's second tile on the ticket is taken at the top of the tile -- measured slower: 14.2 vs 14.0 us per step,
// 16.4 vs 15.5 us alone; profiles/NOTES.md 34)
#ifndef GS_EXPAND_W0_LIGHT
#define GS_EXPAND_W0_LIGHT 1
#endif
#ifndef GS_EXPAND_W0_ROWS
#define GS_EXPAND_W0_ROWS 0
#endif
#ifndef GS_EARLY_TICKET
#define GS_EARLY_TICKET 0
#endif
#ifndef GS_PREP_LATE_TRIGGER
#define GS_PREP_LATE_TRIGGER 1
#endif
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
       uint32_t expand_epoch_wanted, uint32_t blocks_wanted) {
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
#if GS_PREP_LATE_TRIGGER
    if (!overlap_previous)
#endif
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    auto may_write = [&]() {
        if (tid == 0 && (int32_t)(epoch_seen - expand_epoch_wanted) < 0) {
            wait_epoch(fb.expand_epoch, expand_epoch_wanted, fb.fault, kFaultPrepEpoch);
            epoch_seen = expand_epoch_wanted;
        }
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
        if (i < n_pods) r = req[i];
        // With overlapping batches the k_prep of the batch before the previous one may still be polling this
        // buffer set's bucket counts: nothing of the set is touched before the expand kernel that last read it
        // (and therefore that k_prep) is known to be over.  The epoch was read together with the request above.
        if (overlap_previous) may_write();
        {
            // arrival number inside the bucket: the lanes of a warp that share a bucket go to L2 as one atomic
            // (real request mixes have a handful of popular sizes: one address would take 10 000 atomics in turn)
            const u64 bq = r >> shift;
            const uint32_t b = i < n_pods ? (bq < kBuckets - 1 ? (uint32_t)bq : kBuckets - 1) : kBuckets + lane;
            const uint32_t same = __match_any_sync(0xffffffffu, b);
            const uint32_t leader = __ffs(same) - 1;
            uint32_t first = 0;
            if (lane == leader && i < n_pods) first = atomicAdd(&ghist[b], (uint32_t)__popc(same));
            first = __shfl_sync(0xffffffffu, first, leader);
            if (i < n_pods) sl = (b << 16) | (first + __popc(same & ((1u << lane) - 1u)));
        }
#ifdef GS_EXPAND_TRACE
        if (sl == 0xffffffffu) return;  // (never true: makes the stamp below wait for the atomic's result)
        GS_KTRACE(2 * (expand_epoch_wanted + 1) + fb.set_index, 0, 3);
#endif
        // The counts are complete when they add up to the number of pods: every block polls the 512 counts (the
        // pod blocks are the first blocks of a small grid of small blocks: all of them are resident), and the
        // poll that finds them complete already holds the values the scan needs.
        uint4 c4;
        uint32_t run;
        SpinGuard poll_guard;
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
            poll_guard.tick(fb.fault, kFaultPodCounts);
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
        // (a count of finished blocks over all launches on this buffer set, never reset: a CTA of k_expand that is let in
        // after this point finds it at `blocks_wanted`)
        is_last = atomicAdd(fb.prep_done, 1u) + 1u == blocks_wanted;
    }
    __syncthreads();
#if GS_PREP_LATE_TRIGGER
    // (overlapping batches) the expand kernel of this batch is let in only now, block by block: when its first CTA
    // starts, every block of this grid has fenced its results, and that CTA need not wait for an epoch before it asks
    // for its first tile's inputs
    if (overlap_previous) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
    if (is_last) {
        reinterpret_cast<uint4*>(fb.pod_hist)[tid] = make_uint4(0, 0, 0, 0);
        __syncthreads();
        if (tid == 0) {
            __threadfence();
            atomicAdd(fb.prep_epoch, 1u);  // results complete AND bucket counts re-armed
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

// A tile is kTileRows consecutive grouped pods x kTileWords words (4096 nodes); its rows span request
// buckets [bmin, bmax].  A node with q' > bmax passes in every row (base word), one with q' < bmin in
// none; only nodes with bmin <= q' <= bmax need the exact comparison key >= req per row.  Words holding
// such nodes ("partial" words) are settled per tile in one of two ways:
//   sparse tile (<= kSparseCap partial words, each with <= kSparseMax compared nodes): lane = row, exact
//     compares against the few keys, results in a small shared-memory patch;
//   dense tile (anything else): the tile's rows are sorted, every node of a partial word gets
//     cnt = #{rows : req <= key} by binary search (lane = node), and row i of the sorted tile has the
//     node's bit set iff cnt > i; the counts are kept bit-sliced.
// Every other word goes from registers straight to global memory.  Stores are 16 bytes per lane, 512
// contiguous bytes per warp and row.  (Bulk asynchronous stores of 512-byte row segments from a
// shared-memory tile were measured slower than this: profiles/NOTES.md.)
//
// Node keys reach shared memory only through bulk asynchronous copies: an ordinary load issued by an SM
// that is streaming stores waits behind them for microseconds (profiles/NOTES.md), and so does everything
// that depends on it.  The copies are issued as soon as the tile is classified, and the lanes whose four
// words are the same in every row store all their rows while the keys are in flight.
//
// Keys are compared as 32-bit saturated values (node_key32 = min(key, 2^32 - 1)), which is exact for every
// request below 2^32; a tile holding a larger request (a wrapped negative sum, pkg/cache/nodeinfo.go:217)
// is a "wide" tile and takes the dense path over the full 64-bit keys, half a tile of keys at a time.
constexpr int kSparseMax = 4;       // compared nodes per word on the lane = row path
constexpr uint32_t kSparseCap = 16; // partial words per tile on that path
constexpr uint32_t kPatchStride = kTileRows + 1;  // odd: phase D reads one row of several patched words
constexpr uint32_t kStageWords = 4096;            // 16 KB: the tile's 4096 32-bit keys, or 2048 64-bit ones

struct ExpandSmem {
    // dense tile: node keys; sparse tile: [kSparseCap][32] keys, then [kSparseCap][kPatchStride] patch
    uint32_t stage[kStageWords];
    u64 req[kTileRows];                 // requests of the tile's rows (~0 beyond the last row)
    u64 req_sorted[kTileRows];          // the same in ascending order (dense tiles)
    __align__(16) uint32_t req_sorted32[kTileRows];  // ... saturated to 32 bits (exact in a tile that is not "wide")
    __align__(16) uint32_t pivot[kTileRows / 4];     // req_sorted32[3], [7], [11], ...: the search's upper levels
    u64 off_sorted[kTileRows];
    u64 row_off[kTileRows];             // word offset of every tile row's output row (pod index x row stride)
    uint32_t basew[kTileWords];         // nodes passing in every row
    uint32_t somew[kTileWords];         // nodes that can pass in some row (q' >= bmin)
    uint32_t pmask[kTileWords];         // nodes that need the comparison (some & ~base)
    uint32_t cntp[kTileWords * 8];      // dense tiles: per-node pass counts (0..64), bit-sliced, 7 planes used
    uint8_t plist[kTileWords];          // the partial words
    uint8_t pidx[kTileWords];           // position in plist, 0xff = the word is the same in every row
    uint32_t n_partial, n_heavy, next_tile;
};
constexpr uint32_t kSparseKeyWords = kSparseCap * 32;
static_assert(kSparseKeyWords + kSparseCap * kPatchStride <= kStageWords, "sparse keys + patch share the stage");

// Inputs of one tile, brought in by bulk asynchronous copies (TMA) that complete on an mbarrier, one tile
// ahead.
constexpr int kPrefetch = 2;
struct __align__(128) TileInputs {
    uint32_t planes[kRankPlanes * kTileWords];  // bucket-id planes of the tile's 128 words
    u64 req[kTileRows];                         // requests of the tile's rows
    uint16_t perm[kTileRows];                   // their pod indices
    uint16_t bucket[kTileRows];                 // their 1-based buckets
};
constexpr uint32_t kTileInputBytes = sizeof(TileInputs);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
// try_wait suspends the thread for a while by itself; the deadline is only looked at when it keeps failing
__device__ __noinline__ void mbar_wait_slow(uint64_t* bar, uint32_t parity, unsigned int* fault, unsigned int where) {
    SpinGuard g;
    while (!mbar_try_wait(bar, parity)) g.tick(fault, where);
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity, unsigned int* fault, unsigned int where) {
    if (!mbar_try_wait(bar, parity)) mbar_wait_slow(bar, parity, fault, where);
}
__device__ __forceinline__ void bulk_load(void* sdst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(sdst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// orders this thread's earlier generic-proxy accesses to shared memory before its later asynchronous copies
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

#ifdef GS_EXPAND_TRACE
// development builds only: per CTA and tile, SM cycle stamps of the phases (tools/expand_trace.py)
constexpr int kTraceTiles = 6, kTraceSlots = 16;
__device__ unsigned long long g_trace[148 * 5 * kTraceTiles * kTraceSlots];
#define GS_TRACE(slot) \
    do { if (tid == 0 && it < kTraceTiles) g_trace[((size_t)blockIdx.x * kTraceTiles + it) * kTraceSlots + (slot)] = clock64(); } while (0)
#define GS_TRACE_VAL(slot, v) \
    do { if (tid == 0 && it < kTraceTiles) g_trace[((size_t)blockIdx.x * kTraceTiles + it) * kTraceSlots + (slot)] = (v); } while (0)
#else
#define GS_TRACE(slot) do {} while (0)
#define GS_TRACE_VAL(slot, v) do {} while (0)
#endif

template <bool kVec>
__global__ void __launch_bounds__(kExpandThreads, kExpandThreads == 256 ? 5 : 8)
k_expand(FilterBuffers fb, uint32_t n_pods, uint32_t words_out, uint32_t row_stride, uint32_t* __restrict__ out,
         uint32_t n_col_blocks, uint32_t inv_col_blocks, uint32_t n_tiles, uint32_t overlap_previous,
         uint32_t prep_epoch_wanted, uint32_t prep_blocks_wanted) {
    __shared__ __align__(128) ExpandSmem sm;
    const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr uint32_t kWarps = kExpandThreads / 32;
    static_assert(kExpandThreads >= kTileWords && kExpandThreads >= kTileRows, "phase A: one thread per word, one per row");
    const uint32_t* __restrict__ planes = fb.planes;

    __shared__ TileInputs tin[kPrefetch];
    __shared__ __align__(8) uint64_t tin_bar[kPrefetch];
    __shared__ __align__(8) uint64_t key_bar;  // completion of the key copies of the current tile
    uint32_t key_phase = 0;

    // thread 0: request the inputs of the tile with ticket `ticket_no` into stage `stage`
    auto request = [&](uint32_t ticket_no, uint32_t stage) {
        if (ticket_no >= n_tiles) return;
        const uint32_t t = n_tiles - 1u - ticket_no;  // high buckets first: that is where node keys cluster
        const uint32_t rt = inv_col_blocks ? __umulhi(t, inv_col_blocks) : t, cb = t - rt * n_col_blocks;  // t / n_col_blocks
        const uint32_t j0 = rt * kTileRows;
        TileInputs& in = tin[stage];
        mbar_expect_tx(&tin_bar[stage], kTileInputBytes);
        bulk_load(in.planes, planes + (size_t)cb * kRankPlanes * kTileWords, sizeof(in.planes), &tin_bar[stage]);
        bulk_load(in.req, fb.grouped_req + j0, sizeof(in.req), &tin_bar[stage]);
        bulk_load(in.perm, fb.perm + j0, sizeof(in.perm), &tin_bar[stage]);
        bulk_load(in.bucket, fb.row_bucket + j0, sizeof(in.bucket), &tin_bar[stage]);
    };

    // Tile schedule: the first tile of a CTA is blockIdx.x; every later one is a ticket taken during the
    // current tile (after its dense work, if any), so a CTA never holds more than one tile ahead and
    // CTAs stuck in a dense tile simply take fewer.  (Tickets taken two tiles ahead hide the ticket's round
    // trip but hoard: an isolated launch went from 17.6 to 20.2 us; profiles/NOTES.md.)
    uint32_t ticket = 0;  // thread 0
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < kPrefetch; ++i) mbar_init(&tin_bar[i], 1);
        mbar_init(&key_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        sm.n_partial = 0;
        sm.n_heavy = 0;
    }
    // Everything above overlapped the tail of k_prep; its results are needed from here on.  Normally this
    // grid waits for all earlier work in the stream.  When the host knows that the work before this batch's
    // k_prep is the expand kernel of another batch (other work buffers, other output), this grid only waits for
    // its own k_prep, by epoch, and so fills the SM slots that the previous expand kernel frees one by one
    // instead of idling through its tail.  A following grid is let in after that point.
    GS_KTRACE(2 * prep_epoch_wanted + fb.set_index, 1, 0);
    bool first_requested = false;
    if (!overlap_previous) {
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    } else {
        // the next batch's k_prep may start as soon as this grid is resident: it shares nothing with this batch's
        // k_prep, which may still be running (each buffer set has its own bucket counts)
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#if GS_PREP_LATE_TRIGGER
        // This grid is k_prep's programmatic dependent and k_prep's blocks trigger only after they have fenced what
        // they wrote: a CTA that runs here runs after all of them.  The first tile's inputs are therefore asked for at
        // once, and the count of k_prep's finished blocks -- read at the same time, not before -- only confirms it.  Should it ever not (a launch
        // order this file does not produce), the CTA does what the other build does: waits for the epoch and asks
        // again; the count of such CTAs is readable through gs_debug_early_starts().
        if (tid == 0) {
            asm volatile("fence.proxy.async;" ::: "memory");
            request(blockIdx.x, 0);
            uint32_t seen;  // (relaxed: it confirms, it does not order; the path that waits uses an acquiring read)
            asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(seen) : "l"(fb.prep_done) : "memory");
#ifdef GS_TEST_FORCE_EARLY_START
            if ((blockIdx.x & 7u) == 3u) seen = prep_blocks_wanted - 1u;  // (test builds: exercise the path)
#endif
            if ((int32_t)(seen - prep_blocks_wanted) < 0) {
                atomicAdd(fb.early_starts, 1u);
                if (blockIdx.x < n_tiles) mbar_wait(&tin_bar[0], 0, fb.fault, kFaultTileInputs);  // the copy asked for above
                wait_epoch(fb.prep_epoch, prep_epoch_wanted, fb.fault, kFaultExpandEpoch);
                asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(&tin_bar[0])) : "memory");
                mbar_init(&tin_bar[0], 1);
                asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
                asm volatile("fence.proxy.async;" ::: "memory");
                request(blockIdx.x, 0);
            }
        }
        first_requested = true;
#else
        if (tid == 0) {
            wait_epoch(fb.prep_epoch, prep_epoch_wanted, fb.fault, kFaultExpandEpoch);
            asm volatile("fence.proxy.async;" ::: "memory");  // the bulk copies below read what k_prep wrote
        }
        __syncthreads();
#endif
    }
    GS_KTRACE(2 * prep_epoch_wanted + fb.set_index, 1, 1);
    if (tid == 0 && !first_requested) request(blockIdx.x, 0);
    __syncthreads();
    uint32_t cur = blockIdx.x;

    for (uint32_t it = 0; cur < n_tiles; ++it) {
        const uint32_t stage = it % kPrefetch, parity = (it / kPrefetch) & 1u;
        const uint32_t tile = n_tiles - 1u - cur;
        const uint32_t rt = inv_col_blocks ? __umulhi(tile, inv_col_blocks) : tile, cb = tile - rt * n_col_blocks;
        const uint32_t j0 = rt * kTileRows;
        const uint32_t rows = min((uint32_t)kTileRows, n_pods - j0);
        const uint32_t w0 = cb * kTileWords;
        const TileInputs& in = tin[stage];
        GS_TRACE(0);
#if GS_EARLY_TICKET
        // From a CTA's second tile on the ticket for the tile after this one is taken right here, so that its round trip
        // is over when it is wanted (just before this tile's stores).  Dense tiles sit at the high buckets, which are
        // the CTAs' FIRST tiles; only there is the ticket taken after the tile is known not to be dense.
        bool have_ticket = false;  // thread 0
        if (tid == 0 && it != 0) {
            ticket = atomicAdd(&fb.sched[0], 1u);
            have_ticket = true;
        }
#else
        const bool have_ticket = false;
#endif
        mbar_wait(&tin_bar[stage], parity, fb.fault, kFaultTileInputs);
        GS_TRACE(1);

        // A: one thread per word (threads 0..127), one per row (threads 128..191)
        uint32_t my_pi = 0xffu;  // threads 0..127: position of the thread's word among the partial ones
        int big_req = 0;         // threads 128..191: the row's request does not fit 32 bits
        if (tid < kTileWords) {
            uint32_t a[kRankPlanes];
#pragma unroll
            for (int k = 0; k < kRankPlanes; ++k) a[k] = in.planes[k * kTileWords + tid];
            const uint32_t bmin = in.bucket[0], bmax = in.bucket[rows - 1];
            // the last bucket is open-ended: no node is known to cover all of it
            const uint32_t t_all = bmax == kBuckets ? kBuckets + 2u : bmax + 1u;
            const uint32_t all_rows = planes_ge(a, t_all);
            const uint32_t some_rows = planes_ge(a, bmin);
            const uint32_t pm = some_rows & ~all_rows;
            sm.basew[tid] = all_rows;
            sm.somew[tid] = some_rows;
            sm.pmask[tid] = pm;
            if (pm) {
                my_pi = atomicAdd(&sm.n_partial, 1u);
                sm.plist[my_pi] = (uint8_t)tid;
                if (__popc(pm) > kSparseMax) sm.n_heavy = 1u;
            }
            sm.pidx[tid] = (uint8_t)my_pi;
        }
        // (a CTA with fewer threads than words + rows lets its first threads take the rows as well)
        constexpr bool kRowThreadsApart = kExpandThreads >= kTileWords + kTileRows;
        if (kRowThreadsApart ? (tid >= kTileWords && tid < kTileWords + kTileRows) : (tid < kTileRows)) {
            const uint32_t r = kRowThreadsApart ? tid - kTileWords : tid;
            const uint32_t row = (uint32_t)in.perm[r];
            const u64 rq = in.req[r];
            sm.row_off[r] = (u64)row * row_stride;
            sm.req[r] = r < rows ? rq : ~0ull;
            big_req = (r < rows && (rq >> 32) != 0) ? 1 : 0;
        }
        // phase A results are visible; nobody reads this stage's inputs any more
        const bool wide = __syncthreads_or(big_req) != 0;

        const uint32_t np = sm.n_partial;
        const bool dense = np != 0 && (sm.n_heavy != 0 || np > kSparseCap || wide);
        const bool sparse = np != 0 && !dense;
        GS_TRACE(2);
        GS_TRACE_VAL(8, (unsigned long long)np | ((unsigned long long)dense << 16) | ((unsigned long long)wide << 17) | ((unsigned long long)tile << 32));
        {
            unsigned long long gt;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(gt));
            GS_TRACE_VAL(9, gt);
        }

        // the keys this tile needs, on their way while the rows that do not need them are stored
        u64* const stage64 = reinterpret_cast<u64*>(sm.stage);
        if (dense) {
            if (tid == 0) {
                fence_async_smem();
                mbar_expect_tx(&key_bar, kStageWords * 4u);
                if (wide)  // the keys of the first 64 words
                    bulk_load(stage64, fb.node_key + (size_t)w0 * 32, kStageWords * 4u, &key_bar);
                else
                    bulk_load(sm.stage, fb.node_key32 + (size_t)w0 * 32, kStageWords * 4u, &key_bar);
            }
        } else if (sparse) {
            // each partial word's thread fetches the 32 keys of its word (128 bytes); the expected byte count
            // may be posted after some of the copies have already landed, which the barrier allows
            if (tid == 0) mbar_expect_tx(&key_bar, np * 128u);
            if (my_pi != 0xffu) {
                fence_async_smem();
                bulk_load(&sm.stage[my_pi * 32], fb.node_key32 + (size_t)(w0 + tid) * 32, 128u, &key_bar);
            }
        }

        if (!dense && tid == 0 && !have_ticket) ticket = atomicAdd(&fb.sched[0], 1u);  // consumed just before the stores

        // Dense tiles want the tile's rows in ascending request order: then a node's column is a prefix of
        // ones and one binary search per node replaces a comparison per row.
        // (rank by counting; the kSortParts lanes of a row each count a share of the other rows)
        constexpr uint32_t kSortParts = kExpandThreads / kTileRows >= 4 ? 4u : 1u;
        if (dense && tid < kTileRows * kSortParts) {
            const uint32_t r = tid / kSortParts, part = tid % kSortParts;
            const u64 mine = sm.req[r];
            uint32_t rank = 0;
#pragma unroll
            for (uint32_t i = 0; i < kTileRows / kSortParts; ++i) {
                const uint32_t o = part * (kTileRows / kSortParts) + i;
                const u64 other = sm.req[o];
                rank += (other < mine || (other == mine && o < r)) ? 1u : 0u;
            }
            if (kSortParts == 4) {
                rank += __shfl_xor_sync(0xffffffffu, rank, 1);
                rank += __shfl_xor_sync(0xffffffffu, rank, 2);
            }
            if (part == 0) {
                sm.req_sorted[rank] = mine;
                const uint32_t mine32 = (mine >> 32) != 0 ? 0xffffffffu : (uint32_t)mine;
                sm.req_sorted32[rank] = mine32;
                if ((rank & 3u) == 3u) sm.pivot[rank >> 2] = mine32;
                sm.off_sorted[rank] = sm.row_off[r];
            }
        }
        GS_TRACE(3);

        if (!dense && tid == 0) {  // the next tile is known: start fetching its inputs
            sm.next_tile = gridDim.x + ticket;
            request(sm.next_tile, stage ^ 1u);
        }
        // D, first part: lane owns 4 words, warp owns rows warp, warp + kWarps, ...; 16 bytes per lane, 512
        // contiguous bytes per warp and row.  Lanes whose four words are the same in every row store now.
        // In a narrow last column block the lanes of a warp are spread over several rows instead (lanes_per_row
        // lanes each), so that a store instruction still covers 512 bytes: a warp's worth of 128-byte row ends
        // costs as much store-pipe time as full rows.
        const uint32_t width4 = min(32u, (words_out - w0) >> 2);  // 16-byte columns of this tile (kVec)
        const uint32_t lane_shift = width4 > 16 ? 5u : width4 > 8 ? 4u : width4 > 4 ? 3u : width4 > 2 ? 2u : width4 > 1 ? 1u : 0u;
        const uint32_t lcol = lane & ((1u << lane_shift) - 1u);      // the lane's 16-byte column
        const uint32_t row0 = (warp << (5u - lane_shift)) + (lane >> lane_shift);  // its first row
        const uint32_t row_step = kWarps << (5u - lane_shift);
        const uint32_t wcol = w0 + lcol * 4;
        uint4 b4 = make_uint4(0, 0, 0, 0);
        uint32_t p4 = 0xffffffffu;
        if (kVec) {
            b4 = *reinterpret_cast<const uint4*>(&sm.basew[lcol * 4]);
            p4 = *reinterpret_cast<const uint32_t*>(&sm.pidx[lcol * 4]);
            if (wcol < words_out && p4 == 0xffffffffu) {
                uint32_t* dst = out + wcol;
#if GS_EXPAND_W0_LIGHT
                // Warp 0 is the one that waits for the ticket and issues the next tile's input copies before it stores,
                // and the tile ends when its last warp has stored: in a full-width tile warp 0 keeps only the first
                // kW0Rows of its eight rows (0, 8, 16, ...), the others go round warps 1..7, one row each.
                constexpr uint32_t kW0Rows = GS_EXPAND_W0_ROWS;
                const bool light = lane_shift == 5u;
                const uint32_t my_end = (light && warp == 0) ? min(rows, kW0Rows * kWarps) : rows;
#pragma unroll 4
                for (uint32_t rr = row0; rr < my_end; rr += row_step)
                    __stcs(reinterpret_cast<uint4*>(dst + sm.row_off[rr]), b4);
                if (light && warp != 0) {
#pragma unroll
                    for (uint32_t i = kW0Rows; i < kTileRows / kWarps; ++i) {
                        const uint32_t rr = i * kWarps;  // a row of warp 0
                        if ((i - kW0Rows) % (kWarps - 1) + 1 == warp && rr < rows)
                            __stcs(reinterpret_cast<uint4*>(dst + sm.row_off[rr]), b4);
                    }
                }
#else
#pragma unroll 4
                for (uint32_t rr = row0; rr < rows; rr += row_step)
                    __stcs(reinterpret_cast<uint4*>(dst + sm.row_off[rr]), b4);
#endif
            }
        } else {
            for (uint32_t rr = warp; rr < rows; rr += kWarps) {
                uint32_t* dst = out + sm.row_off[rr] + w0;
#pragma unroll
                for (int c = 0; c < kTileWords / 32; ++c) {
                    const uint32_t wl = c * 32 + lane;
                    if (w0 + wl < words_out && sm.pidx[wl] == 0xffu) __stcs(dst + wl, sm.basew[wl]);
                }
            }
        }

        if (np) {  // uniform
            GS_TRACE(4);
            if (dense) {
                // per node of every partial word: cnt = #{rows : req <= key}, kept bit-sliced (plane-major)
                __syncthreads();  // req_sorted is complete
                GS_TRACE(10);
                const uint32_t passes = wide ? 2u : 1u;
                for (uint32_t pass = 0; pass < passes; ++pass) {
                    if (pass && tid == 0) {  // wide tile: the keys of words 64..127 replace those of 0..63
                        fence_async_smem();
                        mbar_expect_tx(&key_bar, kStageWords * 4u);
                        bulk_load(stage64, fb.node_key + (size_t)(w0 + 64u) * 32, kStageWords * 4u, &key_bar);
                    }
                    mbar_wait(&key_bar, key_phase, fb.fault, kFaultTileKeys);
                    key_phase ^= 1u;
                    GS_TRACE(11);
                    // Each warp owns partial words warp, warp + 8, ... and runs several of them at a time, so that
                    // their independent searches share each round trip to shared memory (on an SM that is streaming
                    // stores a dependent shared-memory access costs the better part of a thousand cycles).  With
                    // 32-bit operands: eight words at a time, the 64 sorted requests searched as 4 blocks of 4
                    // groups of 4 -- one comparison round against the block ends (registers), one against the
                    // chosen block's group ends and one against the chosen group: two 16-byte loads per word
                    // instead of six dependent ones.
                    if (!wide) {
                        constexpr int kWay = 8;
                        static_assert(kTileRows == 64, "three levels of four");
                        const uint4 ends = make_uint4(sm.pivot[3], sm.pivot[7], sm.pivot[11], 0u);  // (the last block has no end to test)
                        for (uint32_t k0 = warp; k0 < np; k0 += kWay * kWarps) {
                            uint32_t cnt[kWay], key[kWay];
#pragma unroll
                            for (int u = 0; u < kWay; ++u) {
                                const uint32_t k = k0 + u * kWarps;
                                key[u] = k < np ? sm.stage[(uint32_t)sm.plist[k] * 32 + lane] : 0u;
                                cnt[u] = 16u * ((ends.x <= key[u]) + (ends.y <= key[u]) + (ends.z <= key[u]));  // 0..48
                            }
#pragma unroll
                            for (int u = 0; u < kWay; ++u) {
                                const uint4 gq = *reinterpret_cast<const uint4*>(&sm.pivot[cnt[u] >> 2]);
                                cnt[u] += 4u * ((gq.x <= key[u]) + (gq.y <= key[u]) + (gq.z <= key[u]));
                            }
#pragma unroll
                            for (int u = 0; u < kWay; ++u) {
                                const uint4 e = *reinterpret_cast<const uint4*>(&sm.req_sorted32[cnt[u]]);
                                cnt[u] += (e.x <= key[u]) + (e.y <= key[u]) + (e.z <= key[u]) + (e.w <= key[u]);
                            }
#pragma unroll
                            for (int u = 0; u < kWay; ++u) {
                                const uint32_t k = k0 + u * kWarps;
                                if (k >= np) continue;  // warp-uniform
                                const uint32_t t = sm.plist[k];
                                uint32_t c = min(cnt[u], rows);  // a saturated key also matches the padding rows
                                if (!((sm.somew[t] >> lane) & 1u)) c = 0;  // (see below)
                                uint32_t mine = 0;
#pragma unroll
                                for (int q = 0; q < kDensePlanes; ++q) {
                                    const uint32_t bal = __ballot_sync(0xffffffffu, (c >> q) & 1u);
                                    if ((int)lane == q) mine = bal;
                                }
                                if (lane < kDensePlanes) sm.cntp[lane * kTileWords + t] = mine;
                            }
                        }
                    }
                    // ... with the 64-bit operands of a wide tile: four at a time, plain binary search
                    for (uint32_t k0 = warp; wide && k0 < np; k0 += 4 * kWarps) {
                        uint32_t cnt[4], t[4];
                        u64 key[4];
                        bool on[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const uint32_t k = k0 + u * kWarps;
                            t[u] = k < np ? sm.plist[k] : 0u;
                            on[u] = k < np && (!wide || (t[u] >> 6) == pass);  // warp-uniform
                            key[u] = 0ull;
                            if (on[u]) key[u] = wide ? stage64[(t[u] & 63u) * 32 + lane] : (u64)sm.stage[t[u] * 32 + lane];
                            cnt[u] = 0;
                        }
                        // branch-free upper bound over the 64 sorted requests
#pragma unroll
                        for (uint32_t step = kTileRows / 2; step > 0; step >>= 1) {
#pragma unroll
                            for (int u = 0; u < 4; ++u)
                                if (sm.req_sorted[cnt[u] + step - 1] <= key[u]) cnt[u] += step;
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            if (!on[u]) continue;
                            uint32_t c = cnt[u];
                            if (c == kTileRows - 1 && sm.req_sorted[kTileRows - 1] <= key[u]) c = kTileRows;
                            c = min(c, rows);  // a key of ~0 also matches the padding rows
                            // nodes outside `some` never pass (q' == 0 stores key 0, which would match a zero request)
                            if (!((sm.somew[t[u]] >> lane) & 1u)) c = 0;
                            uint32_t mine = 0;
#pragma unroll
                            for (int q = 0; q < kDensePlanes; ++q) {
                                const uint32_t bal = __ballot_sync(0xffffffffu, (c >> q) & 1u);
                                if ((int)lane == q) mine = bal;
                            }
                            if (lane < kDensePlanes) sm.cntp[lane * kTileWords + t[u]] = mine;
                        }
                    }
                    GS_TRACE(12);
                    // the ticket of a dense tile is taken late (see the schedule note above): when this warp's
                    // share of the last pass is done, so that the round trip overlaps the wait for the other warps
                    if (tid == 0 && pass + 1 == passes && !have_ticket) ticket = atomicAdd(&fb.sched[0], 1u);
                    __syncthreads();  // counts complete; the key area may be overwritten
                    GS_TRACE(13);
                }
                if (tid == 0) {
                    sm.next_tile = gridDim.x + ticket;
                    request(sm.next_tile, stage ^ 1u);
                }
            } else {
                // C: lane = row; the compared nodes' keys come from the staged 32 keys of the word
                mbar_wait(&key_bar, key_phase, fb.fault, kFaultTileKeys);
                key_phase ^= 1u;
                uint32_t* const patch = sm.stage + kSparseKeyWords;
                for (uint32_t k = warp; k < np; k += kWarps) {
                    const uint32_t t = sm.plist[k];
                    const uint32_t base = sm.basew[t];
                    uint32_t pm = sm.pmask[t];
                    u64 kk[kSparseMax];
                    uint32_t kb[kSparseMax];
#pragma unroll
                    for (int i = 0; i < kSparseMax; ++i) {
                        kk[i] = 0;
                        kb[i] = 0;  // unused slots set nothing
                        if (pm) {
                            const uint32_t bit = __ffs(pm) - 1;
                            pm &= pm - 1;
                            kk[i] = sm.stage[k * 32 + bit];
                            kb[i] = 1u << bit;
                        }
                    }
                    for (uint32_t rr = lane; rr < rows; rr += 32) {
                        const u64 rq = sm.req[rr];
                        uint32_t v = base;
#pragma unroll
                        for (int i = 0; i < kSparseMax; ++i)
                            if (kk[i] >= rq) v |= kb[i];
                        patch[k * kPatchStride + rr] = v;
                    }
                }
                __syncthreads();
            }

            // D, second part: the lanes that own a partial word
            GS_TRACE(5);
            const uint32_t* const patch = sm.stage + kSparseKeyWords;
            const u64* __restrict__ toff = dense ? sm.off_sorted : sm.row_off;
            if (kVec) {
                if (wcol < words_out && p4 != 0xffffffffu) {
                    uint32_t* dst = out + wcol;
                    const uint32_t k0 = p4 & 0xffu, k1 = (p4 >> 8) & 0xffu, k2 = (p4 >> 16) & 0xffu, k3 = p4 >> 24;
                    if (dense) {
                        // The lane's count planes do not depend on the row: loaded once, ahead of every store.  Its
                        // rows are row_step (a multiple of 8) apart, so the three low bits of row + 1 are the same
                        // for all of them and the three low planes fold into one value up front.
                        static_assert(kWarps % 8 == 0 && kDensePlanes > 3, "low bits of the lane's rows are fixed");
                        uint4 low = make_uint4(~0u, ~0u, ~0u, ~0u);
#pragma unroll
                        for (int q = 0; q < 3; ++q) {
                            const uint4 pl = *reinterpret_cast<const uint4*>(&sm.cntp[q * kTileWords + lcol * 4]);
                            if (((row0 + 1u) >> q) & 1u) {
                                low.x &= pl.x; low.y &= pl.y; low.z &= pl.z; low.w &= pl.w;
                            } else {
                                low.x |= pl.x; low.y |= pl.y; low.z |= pl.z; low.w |= pl.w;
                            }
                        }
                        uint4 cp[kDensePlanes - 3];
#pragma unroll
                        for (int q = 3; q < kDensePlanes; ++q)
                            cp[q - 3] = *reinterpret_cast<const uint4*>(&sm.cntp[q * kTileWords + lcol * 4]);
                        for (uint32_t rr = row0; rr < rows; rr += row_step) {
                            // bit set iff the node's pass count reaches (sorted) row + 1
                            const uint32_t thr = rr + 1;
                            uint4 ge = low;
#pragma unroll
                            for (int q = 3; q < kDensePlanes; ++q) {
                                const uint4 pl = cp[q - 3];
                                if ((thr >> q) & 1u) {
                                    ge.x &= pl.x; ge.y &= pl.y; ge.z &= pl.z; ge.w &= pl.w;
                                } else {
                                    ge.x |= pl.x; ge.y |= pl.y; ge.z |= pl.z; ge.w |= pl.w;
                                }
                            }
                            uint4 v = b4;
                            if (k0 != 0xffu) v.x = ge.x;
                            if (k1 != 0xffu) v.y = ge.y;
                            if (k2 != 0xffu) v.z = ge.z;
                            if (k3 != 0xffu) v.w = ge.w;
                            __stcs(reinterpret_cast<uint4*>(dst + toff[rr]), v);
                        }
                    } else {
                        for (uint32_t rr = row0; rr < rows; rr += row_step) {
                            uint4 v = b4;
                            if (k0 != 0xffu) v.x = patch[k0 * kPatchStride + rr];
                            if (k1 != 0xffu) v.y = patch[k1 * kPatchStride + rr];
                            if (k2 != 0xffu) v.z = patch[k2 * kPatchStride + rr];
                            if (k3 != 0xffu) v.w = patch[k3 * kPatchStride + rr];
                            __stcs(reinterpret_cast<uint4*>(dst + toff[rr]), v);
                        }
                    }
                }
            } else {
                for (uint32_t rr = warp; rr < rows; rr += kWarps) {
                    uint32_t* dst = out + toff[rr] + w0;
#pragma unroll
                    for (int c = 0; c < kTileWords / 32; ++c) {
                        const uint32_t wl = c * 32 + lane;
                        const uint32_t k = sm.pidx[wl];
                        if (w0 + wl < words_out && k != 0xffu) {
                            uint32_t v;
                            if (dense) {
                                const uint32_t thr = rr + 1;
                                v = ~0u;
#pragma unroll
                                for (int q = 0; q < kDensePlanes; ++q) {
                                    const uint32_t pl = sm.cntp[q * kTileWords + wl];
                                    v = ((thr >> q) & 1u) ? (v & pl) : (v | pl);
                                }
                            } else {
                                v = patch[k * kPatchStride + rr];
                            }
                            __stcs(dst + wl, v);
                        }
                    }
                }
            }
        }
        if (tid == 0) {  // everyone copied these to registers right after phase A
            GS_TRACE(6);
            sm.n_partial = 0;
            sm.n_heavy = 0;
        }
        __syncthreads();  // phase A of the next tile rewrites what this tile was reading
        GS_TRACE(7);
        cur = sm.next_tile;
    }
    GS_KTRACE(2 * prep_epoch_wanted + fb.set_index, 1, 2);
    // the last CTA out re-arms the ticket counter for the next launch and publishes this one
    if (tid == 0) {
        __threadfence();
        if (atomicAdd(&fb.sched[1], 1u) == gridDim.x - 1) {
            fb.sched[0] = 0;
            fb.sched[1] = 0;
            // (the next k_prep on this buffer set waits for the epoch below; by then this batch's k_prep must have
            // re-armed its counters, which it does after letting this grid in)
            if (overlap_previous) wait_epoch(fb.prep_epoch, prep_epoch_wanted, fb.fault, kFaultPrepEpoch);
            __threadfence();
            atomicAdd(fb.expand_epoch, 1u);
            // completion stays transitive along the stream even when this grid did not wait at its start
            // (one lingering thread holds the grid open)
            asm volatile("griddepcontrol.wait;" ::: "memory");
        }
    }
}

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
static bool g_attr_done[64] = {};

static cudaError_t ensure_attrs(int device) {
    if (device >= 0 && device < 64 && g_attr_done[device]) return cudaSuccess;
    cudaError_t e = cudaFuncSetAttribute(k_radix_sort<BindKeyLoader>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kSortSmemBytes);
    if (e != cudaSuccess) return e;
    // k_expand wants as many resident CTAs as its shared memory allows: ask for the largest carve-out
    e = cudaFuncSetAttribute(k_expand<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_expand<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    // static shared memory + the occupancy pad may exceed the 48 KB a kernel gets without asking
    e = cudaFuncSetAttribute(k_expand<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_expand<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
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
                        uint32_t expand_epoch_wanted, uint32_t* blocks_total, cudaStream_t st) {
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
    const uint32_t blocks_wanted = *blocks_total + pod_blocks + node_blocks;  // what the set's block count reaches with this launch
    cudaError_t e;
    if (dev_stride == 8)
        e = cudaLaunchKernelEx(&cfg, k_prep<8>, d_req, n_pods, pod_blocks, shift, fb, d_used, d_cap, d_nodeinfo, n_nodes,
                               dev_stride, n_words, ov, expand_epoch_wanted, blocks_wanted);
    else
        e = cudaLaunchKernelEx(&cfg, k_prep<0>, d_req, n_pods, pod_blocks, shift, fb, d_used, d_cap, d_nodeinfo, n_nodes,
                               dev_stride, n_words, ov, expand_epoch_wanted, blocks_wanted);
    if (e == cudaSuccess) *blocks_total = blocks_wanted;
    return e;
}

cudaError_t launch_expand(int sm_count, const FilterBuffers& fb, uint32_t n_pods, uint32_t words,
                          uint32_t row_stride, uint32_t* d_out, bool overlap_previous, uint32_t prep_epoch_wanted,
                          uint32_t prep_blocks_wanted, cudaStream_t st) {
    if (n_pods == 0 || words == 0) return cudaErrorInvalidValue;  // callers count launches: never a silent no-op
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaError_t ea = ensure_attrs(dev);
        if (ea != cudaSuccess) return ea;
    }
    const bool vec = (row_stride % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_out) & 15u) == 0);
    // 16-byte stores; the row's last 32-byte sector is written whole when the stride leaves room for it (a
    // half-written sector costs a read-modify-write in DRAM: profiles/NOTES.md).  Padding bits are zero.
    uint32_t words_out = words;
    if (vec) {
        words_out = (words + 7u) & ~7u;
        if (words_out > row_stride) words_out = (words + 3u) & ~3u;
    }
    const uint32_t n_col_blocks = (words_out + kTileWords - 1) / kTileWords;
    const uint32_t ov = overlap_previous ? 1u : 0u;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;  // overlap the launch with k_prep's tail
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    const uint32_t n_row_tiles = (n_pods + kTileRows - 1) / kTileRows;
    const uint32_t n_tiles = n_col_blocks * n_row_tiles;
    // Four CTAs per SM stream as fast as five (tools/chunk_write_bw.cu), and the fifth slot's registers and
    // threads are what lets every block of the next batch's k_prep become resident next to a running k_expand
    // (a programmatic dependent is launched only when ALL blocks of its primary have started).  The limit is
    // set through shared memory: static (36 KB) + this unused dynamic pad + 1 KB reserved > 228 KB / 5.
    const uint32_t per_sm = 4;
    constexpr uint32_t kOccupancyPadBytes = 10240;
    uint32_t grid = (uint32_t)sm_count * per_sm;
    if (grid > n_tiles) grid = n_tiles;

    // exact t / n_col_blocks as a multiply-high for every t below 2^32 / n_col_blocks (n_tiles is far
    // smaller); 0 stands for "one column block" (2^32 does not fit)
    const uint32_t inv = n_col_blocks == 1 ? 0u : (uint32_t)((0x100000000ull + n_col_blocks - 1) / n_col_blocks);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kExpandThreads);
    cfg.dynamicSmemBytes = kOccupancyPadBytes;
    cfg.stream = st;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    if (vec)
        return cudaLaunchKernelEx(&cfg, k_expand<true>, fb, n_pods, words_out, row_stride, d_out, n_col_blocks, inv, n_tiles, ov,
                                  prep_epoch_wanted, prep_blocks_wanted);
    return cudaLaunchKernelEx(&cfg, k_expand<false>, fb, n_pods, words_out, row_stride, d_out, n_col_blocks, inv, n_tiles, ov,
                              prep_epoch_wanted, prep_blocks_wanted);
}

}  // namespace gs

#ifdef GS_EXPAND_TRACE
extern "C" int gs_debug_ktrace_reset() {
    static unsigned long long lo[64 * 2 * 6], hi[64 * 2 * 6];
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
extern "C" int gs_debug_expand_trace(unsigned long long* host, size_t n) {
    if (n > sizeof(gs::g_trace) / 8) n = sizeof(gs::g_trace) / 8;
    return (int)cudaMemcpyFromSymbol(host, gs::g_trace, n * 8);
}
#endif
