// K1 (cache rebuild), K2 (incremental pod events) and K4 (ordered per-node binpack).
#include "common.cuh"
#include "kernels.h"
#include "state_device.cuh"

#include <algorithm>

namespace gs {

using u64 = unsigned long long;

// zeroes used[] ahead of a rebuild; k_rebuild is launched as its programmatic dependent and reads its records
// while this runs
__global__ void __launch_bounds__(256) k_zero_u64x2(ulonglong2* d, size_t n2) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += stride) d[i] = make_ulonglong2(0ull, 0ull);
}

__global__ void k_fill_i32(int32_t* d, int32_t v, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) d[i] = v;
}

__global__ void k_validate_records(gs_pod_record* pods, uint32_t first, uint32_t n, const uint32_t* __restrict__ nodeinfo,
                                   uint32_t n_nodes, uint32_t dev_stride) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    gs_pod_record r = load_record(pods + first + i);
    if ((r.flags & GS_POD_VALID) && !legal_device(nodeinfo, n_nodes, dev_stride, r.node_idx, r.dev)) {
        r.flags = 0;
        store_record(pods + first + i, r);
    }
}

// ------------------------------------------------------------------------------------------------
// K1: used[node][dev] = sum of mem over valid, counted records (pkg/cache/deviceinfo.go:41-54 as a
// segmented integer reduction).  Each warp reduces runs of equal (node, dev) with shuffles, so a pod
// table grouped by (node, device) issues one atomic per run; everything else is one 64-bit RED per record.
// ------------------------------------------------------------------------------------------------
constexpr int kRebuildThreads = 256;
constexpr int kRebuildItems = 4;           // records per thread and chunk
constexpr uint32_t kRebuildChunk = kRebuildThreads * kRebuildItems;

// Persistent: a CTA walks chunks blockIdx.x, blockIdx.x + gridDim.x, ... and loads the next chunk's records before it
// works on the current one, so record loads and atomics of different chunks overlap all the way through.
// (A shared-memory window of privatised bins per CTA, which round 1 had, lost to plain global atomics for every table
// order measured -- random slots, slots grouped by node, slots sorted by (node, device): 64-bit shared-memory
// atomics are compare-and-swap loops -- and is gone; profiles/NOTES.md.)
__global__ void __launch_bounds__(kRebuildThreads)
k_rebuild(const gs_pod_record* __restrict__ pods, uint32_t n_slots, uint32_t dev_stride, u64* __restrict__ used) {
    const uint32_t tid = threadIdx.x, lane = tid & 31;
    const uint32_t n_chunks = (n_slots + kRebuildChunk - 1) / kRebuildChunk;

    auto load_chunk = [&](uint32_t chunk, gs_pod_record (&r)[kRebuildItems]) {
#pragma unroll
        for (int it = 0; it < kRebuildItems; ++it) {
            const uint32_t i = chunk * kRebuildChunk + it * kRebuildThreads + tid;
            r[it].flags = 0;
            if (chunk < n_chunks && i < n_slots) r[it] = load_record(pods + i);
        }
    };
    gs_pod_record recs[kRebuildItems], next[kRebuildItems];
    load_chunk(blockIdx.x, recs);
    // used[] is being zeroed by the kernel before this one: nothing is added to it before that is complete
    asm volatile("griddepcontrol.wait;" ::: "memory");
    for (uint32_t chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
        load_chunk(chunk + gridDim.x, next);
#pragma unroll
        for (int it = 0; it < kRebuildItems; ++it) {
            const gs_pod_record r = recs[it];
            uint32_t bin = 0xffffffffu;
            u64 val = 0;
            if ((r.flags & (GS_POD_VALID | GS_POD_COUNTED)) == (GS_POD_VALID | GS_POD_COUNTED)) {
                bin = r.node_idx * dev_stride + (uint32_t)r.dev;
                val = r.mem;
            }
            // segmented inclusive scan over runs of equal bins: one atomic per run and warp
            const uint32_t prev = __shfl_up_sync(0xffffffffu, bin, 1);
            const uint32_t heads = __ballot_sync(0xffffffffu, (lane == 0) || (prev != bin));
            if (heads != 0xffffffffu) {  // warp-uniform: a table in no particular order has no runs, and pays nothing for them
                const uint32_t seg_start = 31u - __clz(heads & (0xffffffffu >> (31 - lane)));
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    u64 t = __shfl_up_sync(0xffffffffu, val, o);
                    if (lane >= seg_start + o) val += t;
                }
            }
            const bool tail = (lane == 31) || ((heads >> (lane + 1)) & 1u);
            if (tail && bin != 0xffffffffu && val != 0) atomicAdd(&used[bin], val);
        }
#pragma unroll
        for (int it = 0; it < kRebuildItems; ++it) recs[it] = next[it];
    }
}

// ------------------------------------------------------------------------------------------------
// K2: pod events (pkg/cache/cache.go:89-127, deviceinfo.go:56-80).  A slot's final state is that of
// its last effective event in the batch and additions commute mod 2^64, so: pass 1 finds the last
// effective event per slot, pass 2 applies exactly those.
// ------------------------------------------------------------------------------------------------
__global__ void k_events_mark(const gs_pod_event* __restrict__ ev, uint32_t n, uint32_t max_pods, int32_t* last_ev,
                              const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes, uint32_t dev_stride) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) event_mark(i, ev, max_pods, last_ev, nodeinfo, n_nodes, dev_stride);
}

__global__ void k_events_apply(const gs_pod_event* __restrict__ ev, uint32_t n, gs_pod_record* pods, uint32_t max_pods,
                               int32_t* last_ev, const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes,
                               uint32_t dev_stride, u64* used) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) event_apply<true>(i, ev, pods, max_pods, last_ev, nodeinfo, n_nodes, dev_stride, used);  // re-arms the scratch
}

// ------------------------------------------------------------------------------------------------
// K4: binds sorted by node (stable, so queue order survives inside a node); the thread that sees the
// first bind of a node walks that node's binds in order: allocateGPUID (nodeinfo.go:209-252) then the
// commit of Allocate (nodeinfo.go:190-203).
// ------------------------------------------------------------------------------------------------
// kRegs: the node's used[] row lives in registers (stride-8 snapshots) for the whole walk and is written
// back once; otherwise the walk reads and writes the row in global memory.
// (sorted_nodes / perm: global memory, or the shared memory of the one-warp kernel below)
template <bool kRegs>
__device__ __forceinline__ void bind_walk(uint32_t j, const gs_bind_req* __restrict__ binds, uint32_t n, const u64* sorted_nodes,
                                          const uint16_t* perm, gs_pod_record* pods, uint32_t max_pods,
                                          const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes, const u64* __restrict__ cap,
                                          uint32_t dev_stride, u64* used, int32_t* __restrict__ out_dev) {
    if (j >= n) return;
    const u64 node64 = sorted_nodes[j];
    if (j > 0 && sorted_nodes[j - 1] == node64) return;  // not the head of its node's run
    const uint32_t node = (uint32_t)node64;
    const bool known = node < n_nodes && nodeinfo_class(nodeinfo[node]) != GS_NODE_NOT_FOUND;
    const uint32_t ndev = known ? nodeinfo_ndev(nodeinfo[node]) : 0u;
    const u64 c = known ? cap[node] : 0ull;
    u64* row = used + (size_t)node * dev_stride;
    u64 reg[8];
    if (kRegs && known) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            ulonglong2 v = reinterpret_cast<const ulonglong2*>(row)[q];
            reg[2 * q] = v.x;
            reg[2 * q + 1] = v.y;
        }
    }
    bool dirty = false;
    for (uint32_t k = j; k < n && sorted_nodes[k] == node64; ++k) {
        const uint32_t bi = perm[k];
        const gs_bind_req b = binds[bi];
        if (!known) {
            out_dev[bi] = GS_BIND_NODE_NOT_FOUND;
            continue;
        }
        int cand = -1;
        u64 cand_mem = 0;
        if (b.req > 0) {
            if (kRegs) {
#pragma unroll
                for (int d = 0; d < 8; ++d) {
                    const u64 avail = c - reg[d];
                    if ((uint32_t)d < ndev && avail >= b.req && (cand == -1 || cand_mem > avail)) {
                        cand = d;
                        cand_mem = avail;
                    }
                }
            } else {
                for (uint32_t d = 0; d < ndev; ++d) {
                    const u64 avail = c - row[d];
                    if (avail >= b.req && (cand == -1 || cand_mem > avail)) {
                        cand = (int)d;
                        cand_mem = avail;
                    }
                }
            }
        }
        gs_pod_record old;
        old.flags = 0;
        if (cand >= 0 && b.slot < max_pods) {
            old = load_record(pods + b.slot);
            // A slot that holds an entry of ANOTHER node cannot be taken over here: that node's used[] row may be in
            // the registers of the thread walking it, and a subtraction from here would be lost.  Refused, untouched.
            if ((old.flags & GS_POD_VALID) && old.node_idx != node) cand = GS_BIND_SLOT_BUSY;
        }
        out_dev[bi] = cand;
        if (cand >= 0 && b.slot < max_pods) {
            if ((old.flags & (GS_POD_VALID | GS_POD_COUNTED)) == (GS_POD_VALID | GS_POD_COUNTED) && old.mem) {
                // the slot already held an entry (same UID bound again): replace, do not double count
                if (kRegs) {
#pragma unroll
                    for (int d = 0; d < 8; ++d)
                        if (d == old.dev) reg[d] -= old.mem;
                    dirty = true;
                } else {
                    row[old.dev] -= old.mem;
                }
            }
            gs_pod_record now;
            now.mem = b.commit_mem;
            now.node_idx = node;
            now.dev = (int16_t)cand;
            now.flags = (uint8_t)(GS_POD_VALID | (b.counted ? GS_POD_COUNTED : 0u));
            now.pad_ = 0;
            store_record(pods + b.slot, now);
            if (b.counted) {
                if (kRegs) {
#pragma unroll
                    for (int d = 0; d < 8; ++d)
                        if (d == cand) reg[d] += b.commit_mem;
                    dirty = true;
                } else {
                    row[cand] += b.commit_mem;
                }
            }
        }
    }
    if (kRegs && dirty) {
#pragma unroll
        for (int q = 0; q < 4; ++q) reinterpret_cast<ulonglong2*>(row)[q] = make_ulonglong2(reg[2 * q], reg[2 * q + 1]);
    }
}

template <bool kRegs>
__global__ void k_bind_scan(const gs_bind_req* __restrict__ binds, uint32_t n, const u64* __restrict__ sorted_nodes,
                            const uint16_t* __restrict__ perm, gs_pod_record* pods, uint32_t max_pods,
                            const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes, const u64* __restrict__ cap,
                            uint32_t dev_stride, u64* used, int32_t* __restrict__ out_dev) {
    bind_walk<kRegs>(blockIdx.x * blockDim.x + threadIdx.x, binds, n, sorted_nodes, perm, pods, max_pods, nodeinfo, n_nodes, cap,
                     dev_stride, used, out_dev);
}

// A queue of at most 32 binds (the extender binds one pod per request) in ONE launch of one warp: every lane counts the
// binds that go before its own (by node, then by position: the stable order the radix sort gives), the order goes to
// shared memory, and the lanes that head a node's run walk it.
template <bool kRegs>
__global__ void __launch_bounds__(32)
k_bind_small(const gs_bind_req* __restrict__ binds, uint32_t n, gs_pod_record* pods, uint32_t max_pods,
             const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes, const u64* __restrict__ cap, uint32_t dev_stride, u64* used,
             int32_t* __restrict__ out_dev) {
    __shared__ u64 s_nodes[32];
    __shared__ uint16_t s_perm[32];
    const uint32_t lane = threadIdx.x;
    const u64 key = lane < n ? (u64)binds[lane].node_idx : ~0ull;
    uint32_t before = 0;
#pragma unroll
    for (uint32_t j = 0; j < 32; ++j) {
        const u64 other = __shfl_sync(0xffffffffu, key, j);
        before += (j < n && (other < key || (other == key && j < lane))) ? 1u : 0u;
    }
    if (lane < n) {
        s_nodes[before] = key;
        s_perm[before] = (uint16_t)lane;
    }
    __syncwarp();
    bind_walk<kRegs>(lane, binds, n, s_nodes, s_perm, pods, max_pods, nodeinfo, n_nodes, cap, dev_stride, used, out_dev);
}

// ------------------------------------------------------------------------------------------------
cudaError_t launch_fill_i32(int32_t* d, int32_t v, size_t n, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    k_fill_i32<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(d, v, n);
    return cudaGetLastError();
}

cudaError_t launch_validate_records(gs_pod_record* d_pods, uint32_t first, uint32_t n, const uint32_t* d_nodeinfo,
                                    uint32_t n_nodes, uint32_t dev_stride, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    k_validate_records<<<(n + 255) / 256, 256, 0, st>>>(d_pods, first, n, d_nodeinfo, n_nodes, dev_stride);
    return cudaGetLastError();
}

cudaError_t launch_rebuild(const gs_pod_record* d_pods, uint32_t n_slots, uint32_t n_node_rows, uint32_t dev_stride,
                           u64* d_used, int sm_count, cudaStream_t st) {
    // used[] has an even number of bytes / 16 only when rows * stride is even; the odd tail word is one more store
    const size_t n_words = (size_t)n_node_rows * dev_stride;
    if (n_words == 0) return cudaSuccess;
    const size_t n2 = n_words / 2;
    if (n_words & 1) {
        cudaError_t e = cudaMemsetAsync(d_used + n_words - 1, 0, sizeof(u64), st);
        if (e != cudaSuccess) return e;
    }
    if (n2) {
        const unsigned blocks = (unsigned)std::min<size_t>((n2 + 255) / 256, (size_t)(sm_count > 0 ? sm_count : 148) * 8);
        k_zero_u64x2<<<blocks, 256, 0, st>>>(reinterpret_cast<ulonglong2*>(d_used), n2);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    if (n_slots == 0) return cudaSuccess;
    const uint32_t n_chunks = (n_slots + kRebuildChunk - 1) / kRebuildChunk;
    const uint32_t resident = (uint32_t)(sm_count > 0 ? sm_count : 148) * 4;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(n_chunks < resident ? n_chunks : resident);
    cfg.blockDim = dim3(kRebuildThreads);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, k_rebuild, d_pods, n_slots, dev_stride, d_used);
}

cudaError_t launch_apply_events(const gs_pod_event* d_events, uint32_t n, gs_pod_record* d_pods, uint32_t max_pods,
                                int32_t* d_last_ev, const uint32_t* d_nodeinfo, uint32_t n_nodes,
                                uint32_t dev_stride, u64* d_used, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    const uint32_t blocks = (n + 255) / 256;
    k_events_mark<<<blocks, 256, 0, st>>>(d_events, n, max_pods, d_last_ev, d_nodeinfo, n_nodes, dev_stride);
    k_events_apply<<<blocks, 256, 0, st>>>(d_events, n, d_pods, max_pods, d_last_ev, d_nodeinfo, n_nodes, dev_stride,
                                           d_used);
    return cudaGetLastError();
}

cudaError_t launch_bind_small(const gs_bind_req* d_binds, uint32_t n, gs_pod_record* d_pods, uint32_t max_pods,
                              const uint32_t* d_nodeinfo, uint32_t n_nodes, const u64* d_cap, uint32_t dev_stride, u64* d_used,
                              int32_t* d_out_dev, cudaStream_t st) {
    if (n == 0 || n > 32) return cudaErrorInvalidValue;
    if (dev_stride == 8)
        k_bind_small<true><<<1, 32, 0, st>>>(d_binds, n, d_pods, max_pods, d_nodeinfo, n_nodes, d_cap, dev_stride, d_used, d_out_dev);
    else
        k_bind_small<false><<<1, 32, 0, st>>>(d_binds, n, d_pods, max_pods, d_nodeinfo, n_nodes, d_cap, dev_stride, d_used, d_out_dev);
    return cudaGetLastError();
}

cudaError_t launch_bind_scan(const gs_bind_req* d_binds, uint32_t n, const u64* d_sorted_nodes, const uint16_t* d_perm,
                             gs_pod_record* d_pods, uint32_t max_pods, const uint32_t* d_nodeinfo, uint32_t n_nodes,
                             const u64* d_cap, uint32_t dev_stride, u64* d_used, int32_t* d_out_dev, cudaStream_t st) {
    if (n == 0) return cudaSuccess;
    if (dev_stride == 8)
        k_bind_scan<true><<<(n + 127) / 128, 128, 0, st>>>(d_binds, n, d_sorted_nodes, d_perm, d_pods, max_pods,
                                                           d_nodeinfo, n_nodes, d_cap, dev_stride, d_used, d_out_dev);
    else
        k_bind_scan<false><<<(n + 127) / 128, 128, 0, st>>>(d_binds, n, d_sorted_nodes, d_perm, d_pods, max_pods,
                                                            d_nodeinfo, n_nodes, d_cap, dev_stride, d_used, d_out_dev);
    return cudaGetLastError();
}

}  // namespace gs
