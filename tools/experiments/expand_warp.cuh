// NOT PART OF THE BUILD.  An experiment of round 2, kept for the record with its measurements in profiles/NOTES.md 21-23: it was
// included from filter_kernels.cu after k_expand and needs what was removed with it -- `planes_low` / `shift` in FilterBuffers
// (bit planes of the key bits below the bucket scale, written by k_prep's node role), kLowPlanes = 10, 256 + 256 words of
// per-group scheduler state per buffer set, and an expand epoch that counts finished warps (commit dc86a05 has all of it).
//
// k_expand_warp: the bitmap writer of K3, one warp per tile, registers only.
//
// A warp tile is kWarpRows consecutive grouped pods x kTileWords words (4096 nodes).  The warp loads what the tile
// needs straight into registers (the ten bucket planes of its column block: one 16-byte load per plane and lane;
// the rows' pod indices, requests and buckets: one row per lane), classifies its 128 words with logic ops, and
// stores the rows from registers: 16 bytes per lane, 512 contiguous bytes per warp and row, evict-first.  There is
// no shared memory, no block-level barrier and no staging: a warp never waits for another warp, so the SM's store
// queue is fed by a few dozen independent warps instead of by CTAs that stop at barriers (the layout's write
// ceiling, measured with tools/warp_tile_bw.cu: 12.5 us alone / 9.3-10 us with consecutive launches overlapped for
// the 62.7 MB of config C4, against 16.1 / 16.0 us for the CTA-tile kernel this one replaces).
//
// Words whose nodes all sit outside the tile's own bucket range are the same in every row (base word).  A tile that
// holds other words (nodes with bmin <= q' <= bmax, which need the exact test key >= req) takes the exact path:
// the tile's rows are sorted by request inside the warp (bitonic, shuffles), every such node gets
// cnt = #{rows : req <= key} by a binary search over the sorted requests (shuffles again), the counts are kept
// bit-sliced per word, and sorted row i of a word is "cnt > i".  Keys are compared as 32-bit saturated values
// (exact for requests below 2^32); a tile holding a larger request takes the same path over the 64-bit keys.
//
// Scheduling: the grid is sized so that every CTA is resident at once (so that the next batch's k_prep, a
// programmatic dependent, can start at once).  The warps form `n_groups` groups (block index modulo n_groups: the
// blocks of a group land on one SM); group c owns the tiles c, c + n_groups, c + 2 n_groups, ...  A warp's first
// tile is fixed by its block index, later ones are tickets from its group's counter, taken at the top of the
// current tile and used after its stores.  One counter per group, because atomics that return a value take turns
// per address (a single counter for all 8 000 tiles of config C4 cost 30 us by itself).  The last warp of a group to
// leave re-arms the group's counters; every warp adds itself to the set's count of finished expand warps, which is
// what the k_prep that next reuses the work buffers waits for.
#pragma once

namespace gs {

#ifdef GS_WARP_ROWS_32
constexpr int kWarpRows = 32;     // rows of a warp tile
constexpr int kCntPlanes = 6;     // bit planes of a pass count 0..kWarpRows
#else
constexpr int kWarpRows = 16;     // rows of a warp tile
constexpr int kCntPlanes = 5;     // bit planes of a pass count 0..kWarpRows
#endif
constexpr uint32_t kSchedGroups = 256;  // ticket counters per buffer set (>= the SM count)

struct RowPair {                  // one row of the tile while it is being sorted
    unsigned long long req;
    uint32_t prm;                 // pod index (padding rows: above every pod index)
};
__device__ __forceinline__ bool row_less(const RowPair& a, const RowPair& b) {
    return a.req < b.req || (a.req == b.req && a.prm < b.prm);
}

// bit b of the result = (count of node b >= t), the counts given as bit planes c[0..kCntPlanes) (LSB first); t per lane
// one step of a bit-sliced ">=" (LSB first): m = all ones where the threshold's bit is set:  m ? (a & ge) : (a | ge).
// Any function of three bits is one LOP3; the table is spelled out so that it stays one.
__device__ __forceinline__ uint32_t ge_step(uint32_t a, uint32_t ge, uint32_t m) {
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xD4;" : "=r"(d) : "r"(a), "r"(ge), "r"(m));
    return d;
}
// bit b of each word = (value of node b >= t), the values given as `n` bit planes of four words (LSB first)
template <int kPlanes>
__device__ __forceinline__ uint4 planes4_ge(const uint4 (&p)[kPlanes], uint32_t t) {
    uint4 ge = make_uint4(~0u, ~0u, ~0u, ~0u);
#pragma unroll
    for (int k = 0; k < kPlanes; ++k) {
        const uint32_t m = 0u - ((t >> k) & 1u);
        ge.x = ge_step(p[k].x, ge.x, m);
        ge.y = ge_step(p[k].y, ge.y, m);
        ge.z = ge_step(p[k].z, ge.z, m);
        ge.w = ge_step(p[k].w, ge.w, m);
    }
    return ge;
}

struct CountPlanes;
__device__ __forceinline__ uint32_t count_ge(const uint32_t* c, uint32_t t) {
    uint32_t ge = 0xffffffffu;
#pragma unroll
    for (int q = 0; q < kCntPlanes; ++q) {
        const uint32_t m = 0u - ((t >> q) & 1u);
        ge = (c[q] & ge & m) | ((c[q] | ge) & ~m);
    }
    return ge;
}

// #{i < rows : sreq[i] <= key} over the kWarpRows sorted requests held one per lane (lanes 0..kWarpRows-1)
template <class T>
__device__ __forceinline__ uint32_t sorted_count(T sreq_mine, T key, uint32_t rows) {
    uint32_t pos = 0;
#pragma unroll
    for (uint32_t step = kWarpRows / 2; step > 0; step >>= 1) {
        const T v = __shfl_sync(0xffffffffu, sreq_mine, pos + step - 1);
        if (v <= key) pos += step;
    }
    const T last = __shfl_sync(0xffffffffu, sreq_mine, kWarpRows - 1);
    if (pos == kWarpRows - 1 && last <= key) pos = kWarpRows;
    return pos < rows ? pos : rows;  // a key of all ones also matches the padding rows
}

struct CountPlanes {  // pass counts (0..kWarpRows) of the 32 nodes of one word, bit-sliced
    uint32_t q[kCntPlanes];
};

// Pass counts of the nodes `pm` of one word.  The word's keys are fetched a chunk at a time with 16-byte loads
// (sixteen saturated 32-bit keys, or eight 64-bit keys), so the loop over the nodes indexes registers.  T = uint32_t:
// every request of the tile is below 2^32 and the saturated keys are exact; T = u64: a tile that holds a larger
// request (a wrapped negative sum of limits, pkg/cache/nodeinfo.go:217).
template <class T>
__device__ __noinline__ CountPlanes word_counts(const T* __restrict__ keys, uint32_t pm, T sreq, uint32_t rows) {
    constexpr int kChunk = 64 / sizeof(T);  // keys per chunk: four 16-byte loads
    CountPlanes c;
#pragma unroll
    for (int q = 0; q < kCntPlanes; ++q) c.q[q] = 0;
    for (int ch = 0; ch < 32 / kChunk; ++ch) {
        const uint32_t pmc = (pm >> (kChunk * ch)) & ((1u << kChunk) - 1u);
        if (!__any_sync(0xffffffffu, pmc != 0)) continue;
        T kk[kChunk];
#pragma unroll
        for (int v = 0; v < kChunk; ++v) kk[v] = 0;
        if (pmc) {
            const uint4* src = reinterpret_cast<const uint4*>(keys + kChunk * ch);
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                const uint4 x = src[v];
                if (sizeof(T) == 4) {
                    kk[4 * v] = (T)x.x; kk[4 * v + 1] = (T)x.y; kk[4 * v + 2] = (T)x.z; kk[4 * v + 3] = (T)x.w;
                } else {
                    kk[2 * v] = (T)(((unsigned long long)x.y << 32) | x.x);
                    kk[2 * v + 1] = (T)(((unsigned long long)x.w << 32) | x.z);
                }
            }
        }
#pragma unroll
        for (int b = 0; b < kChunk; ++b) {
            const bool has = (pmc >> b) & 1u;
            if (!__any_sync(0xffffffffu, has)) continue;
            const uint32_t cnt = sorted_count<T>(sreq, kk[b], rows);
            if (has) {
#pragma unroll
                for (int q = 0; q < kCntPlanes; ++q) c.q[q] |= ((cnt >> q) & 1u) << (kChunk * ch + b);
            }
        }
    }
    return c;
}

template <bool kVec>
__device__ __forceinline__ void store_words(uint32_t* dst, const uint4& v, uint32_t n_valid) {
    if (kVec) {
        __stcs(reinterpret_cast<uint4*>(dst), v);
    } else {
        if (n_valid > 0) __stcs(dst, v.x);
        if (n_valid > 1) __stcs(dst + 1, v.y);
        if (n_valid > 2) __stcs(dst + 2, v.z);
        if (n_valid > 3) __stcs(dst + 3, v.w);
    }
}

constexpr int kExpandWarpThreads = 32;

template <bool kVec>
__global__ void __launch_bounds__(kExpandWarpThreads)
k_expand_warp(FilterBuffers fb, uint32_t n_pods, uint32_t words_out, uint32_t row_stride, uint32_t* __restrict__ out,
              uint32_t n_col_blocks, uint32_t n_tiles, uint32_t n_groups, uint32_t group_warps, uint32_t overlap_previous,
              uint32_t prep_epoch_wanted) {
    const uint32_t lane = threadIdx.x;
    const uint32_t group = blockIdx.x % n_groups;
    uint32_t* const group_ticket = fb.sched + group;
    uint32_t* const group_done = fb.sched + kSchedGroups + group;
    // Everything above this point overlaps the tail of k_prep.  Normally this grid waits for all earlier work in the
    // stream.  When the host knows that the work before this batch's k_prep is the expand kernel of another batch
    // (other work buffers, other output), this grid only waits for its own k_prep, by epoch, and so fills the SM
    // slots that the previous expand kernel frees one by one instead of idling through its tail.
    if (!overlap_previous) {
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    } else {
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        if (lane == 0) wait_epoch(fb.prep_epoch, prep_epoch_wanted, fb.fault, kFaultExpandEpoch);
        __syncwarp();
    }
    uint32_t ticket = blockIdx.x;  // = group + n_groups * (blockIdx.x / n_groups)
    while (ticket < n_tiles) {
        // the next tile's ticket: asked for now, looked at after this tile's stores
        uint32_t next = 0;
        if (lane == 0) next = group + n_groups * (group_warps + atomicAdd(group_ticket, 1u));
        const uint32_t t = n_tiles - 1u - ticket;  // high buckets first: that is where node keys crowd
        const uint32_t rt = t / n_col_blocks, cb = t - rt * n_col_blocks;
        const uint32_t j0 = rt * kWarpRows;
        const uint32_t rows = min((uint32_t)kWarpRows, n_pods - j0);
        const uint32_t w0 = cb * kTileWords;

        // In a narrow last column block the lanes of a warp cover several rows (lanes_per_row lanes each), so that
        // a store instruction still covers 512 bytes.
        uint32_t lane_shift = 5;
        if (kVec) {
            const uint32_t width4 = min(32u, (words_out - w0) >> 2);  // 16-byte columns of this tile
            lane_shift = width4 > 16 ? 5u : width4 > 8 ? 4u : width4 > 4 ? 3u : width4 > 2 ? 2u : width4 > 1 ? 1u : 0u;
        }
        const uint32_t lcol = lane & ((1u << lane_shift) - 1u);  // the lane's 16-byte column
        const uint32_t rsub = lane >> lane_shift;                // which of the rows of one store instruction
        const uint32_t rpi = 32u >> lane_shift;                  // rows per store instruction
        const uint32_t wl = lcol * 4;                            // the lane's first word inside the column block
        const uint32_t wcol = w0 + wl;
        const uint32_t n_valid = wcol < words_out ? min(4u, words_out - wcol) : 0u;

        // tile inputs, all in flight together
        const uint32_t* const plane_src = fb.planes + (size_t)cb * kRankPlanes * kTileWords + wl;
        uint4 all4, some4;
        uint32_t bucket = 0;
        uint4 a0[kRankPlanes];
#pragma unroll
        for (int k = 0; k < kRankPlanes; ++k) a0[k] = *reinterpret_cast<const uint4*>(plane_src + k * kTileWords);
        RowPair mine;
        mine.req = ~0ull;
        mine.prm = 0x10000u + lane;
        if (lane < rows) {
            mine.req = fb.grouped_req[j0 + lane];
            mine.prm = fb.perm[j0 + lane];
            bucket = fb.row_bucket[j0 + lane];
        }
        const uint32_t bmin = __shfl_sync(0xffffffffu, bucket, 0), bmax = __shfl_sync(0xffffffffu, bucket, rows - 1);
        // the last bucket is open-ended: no node is known to cover all of it
        const uint32_t t_all = bmax == kBuckets ? kBuckets + 2u : bmax + 1u;
        all4 = planes4_ge<kRankPlanes>(a0, t_all);
        some4 = planes4_ge<kRankPlanes>(a0, bmin);
        uint32_t pm[4] = {some4.x & ~all4.x, some4.y & ~all4.y, some4.z & ~all4.z, some4.w & ~all4.w};
        if (n_valid == 0) pm[0] = pm[1] = pm[2] = pm[3] = 0;
        const bool partial = __any_sync(0xffffffffu, (pm[0] | pm[1] | pm[2] | pm[3]) != 0);

        uint32_t* const dst = out + wcol;
        // Bucket id and the key bits below the bucket scale, both bit-sliced, ARE the key (for keys below the
        // open-ended last bucket): the exact test is then a bit-sliced comparison against one number per row,
        // no keys fetched, no row order needed.
        const bool planes_exact = partial && bmax < kBuckets && fb.shift <= (uint32_t)kLowPlanes;
        if (!partial) {
            // every word of the tile is the same in all of its rows
#pragma unroll 4
            for (uint32_t r0 = 0; r0 < rows; r0 += rpi) {
                const uint32_t r = r0 + rsub;
                const uint32_t prm = __shfl_sync(0xffffffffu, mine.prm, r & 31u);
                if (r < rows && n_valid) store_words<kVec>(dst + (size_t)prm * row_stride, all4, n_valid);
            }
        } else if (planes_exact) {
            uint4 lo[kLowPlanes];
            {
                const uint32_t* p = fb.planes_low + (size_t)cb * kLowPlanes * kTileWords + wl;
#pragma unroll
                for (int k = 0; k < kLowPlanes; ++k) lo[k] = *reinterpret_cast<const uint4*>(p + k * kTileWords);
            }
            // A node whose bucket id is above the row's passes, one whose id is below fails, one with the row's own id
            // (equal high parts: row buckets below the last one are 1 + (req >> shift), node ids 1 + (key >> shift))
            // passes iff its low bits reach the row's.  Rows come grouped by bucket, so the two id masks change rarely.
            const uint32_t low_mine = (uint32_t)mine.req & ((1u << fb.shift) - 1u);
            uint32_t cur_bucket = 0xffffffffu;
            uint4 above = make_uint4(0, 0, 0, 0), same = above;
            for (uint32_t r0 = 0; r0 < rows; r0 += rpi) {
                const uint32_t r = r0 + rsub;
                const uint32_t prm = __shfl_sync(0xffffffffu, mine.prm, r & 31u);
                const uint32_t b_r = __shfl_sync(0xffffffffu, bucket, r & 31u);
                const uint32_t low_r = __shfl_sync(0xffffffffu, low_mine, r & 31u);
                if (b_r != cur_bucket) {
                    cur_bucket = b_r;
#ifdef GS_EXACT_RELOAD  // (the bucket planes again: a few L1 hits per tile, and forty registers less)
                    uint4 a[kRankPlanes];
#pragma unroll
                    for (int k = 0; k < kRankPlanes; ++k) a[k] = *reinterpret_cast<const uint4*>(plane_src + k * kTileWords);
#else
                    const uint4 (&a)[kRankPlanes] = a0;
#endif
                    above = planes4_ge<kRankPlanes>(a, b_r + 1u);
                    const uint4 from = planes4_ge<kRankPlanes>(a, b_r);
                    same = make_uint4(from.x & ~above.x, from.y & ~above.y, from.z & ~above.z, from.w & ~above.w);
                }
                const uint4 lge = planes4_ge<kLowPlanes>(lo, low_r);  // (planes at and above `shift` are zero on both sides)
                const uint4 v = make_uint4(above.x | (same.x & lge.x), above.y | (same.y & lge.y), above.z | (same.z & lge.z),
                                           above.w | (same.w & lge.w));
                if (r < rows && n_valid) store_words<kVec>(dst + (size_t)prm * row_stride, v, n_valid);
            }
        } else {
            // exact path over the keys themselves (tiles reaching into the open-ended last bucket, or a bucket scale
            // too coarse for the planes).  Sort the tile's rows by request (bitonic over lanes 0..15; the upper lanes
            // hold padding).
            const bool wide = __any_sync(0xffffffffu, lane < rows && (mine.req >> 32) != 0);
#pragma unroll
            for (uint32_t k = 2; k <= (uint32_t)kWarpRows; k <<= 1) {
#pragma unroll
                for (uint32_t j = k >> 1; j > 0; j >>= 1) {
                    RowPair other;
                    other.req = __shfl_xor_sync(0xffffffffu, mine.req, j);
                    other.prm = __shfl_xor_sync(0xffffffffu, mine.prm, j);
                    const bool lower = (lane & j) == 0, asc = (lane & k) == 0 || k == (uint32_t)kWarpRows;
                    const bool take_min = lower == asc;
                    const bool other_less = row_less(other, mine);
                    if (take_min == other_less) mine = other;
                }
            }
            // (lanes 16..31 ran the same exchanges among themselves on padding; nobody reads them)
            CountPlanes c[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
#pragma unroll
                for (int q = 0; q < kCntPlanes; ++q) c[j].q[q] = 0;
                if (!__any_sync(0xffffffffu, pm[j] != 0)) continue;
                const size_t node0 = (size_t)(w0 + wl + j) * 32;
                if (wide)
                    c[j] = word_counts<unsigned long long>(fb.node_key + node0, pm[j], mine.req, rows);
                else
                    c[j] = word_counts<uint32_t>(fb.node_key32 + node0, pm[j], (uint32_t)mine.req, rows);
            }
            const uint32_t allw[4] = {all4.x, all4.y, all4.z, all4.w};
            for (uint32_t r0 = 0; r0 < rows; r0 += rpi) {
                const uint32_t r = r0 + rsub;  // position in sorted order
                const uint32_t prm = __shfl_sync(0xffffffffu, mine.prm, r & 31u);
                uint32_t v[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) v[j] = allw[j] | (count_ge(c[j].q, r + 1u) & pm[j]);
                if (r < rows && n_valid) store_words<kVec>(dst + (size_t)prm * row_stride, make_uint4(v[0], v[1], v[2], v[3]), n_valid);
            }
        }
        ticket = __shfl_sync(0xffffffffu, next, 0);
    }
    // What a finished warp publishes is that it no longer READS the work buffers (its loads have returned: their
    // values went into the stores above), so the k_prep that waits for this count may overwrite them.  The rows it
    // wrote become visible with the kernel's completion, like any kernel's; no fence is needed here -- and a
    // device-scope fence per warp would make every warp wait for the SM's queued stores and drop the SM's L1.
    if (lane == 0) {
        if (atomicAdd(group_done, 1u) == group_warps - 1) {  // last warp of the group: nobody else uses its counters now
            *group_ticket = 0;
            *group_done = 0;
            __threadfence();
        }
        atomicAdd(fb.expand_epoch, 1u);  // finished expand warps on this buffer set, all launches together
    }
    // Completion stays transitive along the stream even when this grid did not wait at its start: no warp leaves
    // before this batch's k_prep is complete (it is: its epoch was seen), and that k_prep's last block in turn
    // stays until the expand kernel before it is complete.
    if (overlap_previous) asm volatile("griddepcontrol.wait;" ::: "memory");
}

}  // namespace gs
