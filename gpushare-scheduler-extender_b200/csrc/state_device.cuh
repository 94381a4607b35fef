// Device functions shared by the pod-event kernels of state_kernels.cu (K2, snapshot-owned scratch) and
// noalloc.cu (K2 with caller-owned buffers and library-resident scratch): one definition of what an event
// means, so the two entry points cannot drift apart.
#pragma once
#include "common.cuh"

namespace gs {

__device__ __forceinline__ gs_pod_record load_record(const gs_pod_record* p) {
    static_assert(sizeof(gs_pod_record) == 16, "record is one 128-bit load");
    uint4 v = *reinterpret_cast<const uint4*>(p);
    gs_pod_record r;
    *reinterpret_cast<uint4*>(&r) = v;
    return r;
}
__device__ __forceinline__ void store_record(gs_pod_record* p, const gs_pod_record& r) {
    *reinterpret_cast<uint4*>(p) = *reinterpret_cast<const uint4*>(&r);
}

// a record may only name a device the node really has: nodeinfo.go:98-108 (`dev, found := n.devs[id]`);
// a device count beyond the row stride of used[] (a caller-built node table can say anything) names nothing
__device__ __forceinline__ bool legal_device(const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes, uint32_t dev_stride,
                                             uint32_t node, int dev) {
    if (node >= n_nodes || dev < 0 || (uint32_t)dev >= dev_stride) return false;
    return (uint32_t)dev < nodeinfo_ndev(nodeinfo[node]);
}

// An event changes the table iff it names a slot of the table and is a REMOVE, or an ADD onto a device
// that exists.  Anything else (unknown op codes included) is skipped, as the reference skips a pod whose
// annotation does not name one of the node's devices.
__device__ __forceinline__ bool event_effective(const gs_pod_event& e, uint32_t n_slots, const uint32_t* __restrict__ nodeinfo,
                                                uint32_t n_nodes, uint32_t dev_stride) {
    if (e.slot >= n_slots) return false;
    if (e.op == GS_EV_REMOVE) return true;
    if (e.op != GS_EV_ADD) return false;
    return legal_device(nodeinfo, n_nodes, dev_stride, e.node_idx, e.dev);
}

// pass 1: remember the last effective event of every slot
__device__ __forceinline__ void event_mark(uint32_t i, const gs_pod_event* __restrict__ ev, uint32_t n_slots, int32_t* last_ev,
                                           const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes, uint32_t dev_stride) {
    const gs_pod_event e = ev[i];
    if (event_effective(e, n_slots, nodeinfo, n_nodes, dev_stride)) atomicMax(&last_ev[e.slot], (int32_t)i);
}

// pass 2: apply exactly those (additions commute mod 2^64, so used[] ends as if every event had been applied
// in array order).  kRearm: leave the slot's scratch entry idle (-1) for the next batch.
template <bool kRearm>
__device__ __forceinline__ void event_apply(uint32_t i, const gs_pod_event* __restrict__ ev, gs_pod_record* pods, uint32_t n_slots,
                                            int32_t* last_ev, const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes,
                                            uint32_t dev_stride, unsigned long long* used) {
    const gs_pod_event e = ev[i];
    if (!event_effective(e, n_slots, nodeinfo, n_nodes, dev_stride)) return;
    if (last_ev[e.slot] != (int32_t)i) return;
    if (kRearm) last_ev[e.slot] = -1;
    const gs_pod_record old = load_record(pods + e.slot);
    if ((old.flags & (GS_POD_VALID | GS_POD_COUNTED)) == (GS_POD_VALID | GS_POD_COUNTED) && old.mem)
        atomicAdd(&used[(size_t)old.node_idx * dev_stride + (uint32_t)old.dev], 0ull - old.mem);
    gs_pod_record now;
    now.mem = 0;
    now.node_idx = 0;
    now.dev = 0;
    now.flags = 0;
    now.pad_ = 0;
    if (e.op == GS_EV_ADD) {
        now.mem = e.mem;
        now.node_idx = e.node_idx;
        now.dev = e.dev;
        now.flags = (uint8_t)(GS_POD_VALID | (e.counted ? GS_POD_COUNTED : 0u));
        if (e.counted && e.mem) atomicAdd(&used[(size_t)e.node_idx * dev_stride + (uint32_t)e.dev], e.mem);
    }
    store_record(pods + e.slot, now);
}

// key of a node = the largest available memory over its devices (nodeinfo.go:254-264, u64 wrap-around
// included); `eligible` = the node can pass at all (class OK and at least one device, nodeinfo.go:123)
__device__ __forceinline__ unsigned long long node_key(const unsigned long long* __restrict__ used, const unsigned long long* __restrict__ cap,
                                                       const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes, uint32_t dev_stride,
                                                       uint32_t node, bool* eligible) {
    *eligible = false;
    if (node >= n_nodes) return 0ull;
    const uint32_t info = nodeinfo[node];
    const uint32_t ndev = nodeinfo_ndev(info);
    if (nodeinfo_class(info) != GS_NODE_OK || ndev == 0) return 0ull;
    *eligible = true;
    const unsigned long long c = cap[node];
    unsigned long long key = 0;
    if (dev_stride == 8) {
        const ulonglong2* row = reinterpret_cast<const ulonglong2*>(used + (size_t)node * 8);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const ulonglong2 u = row[k];
            const unsigned long long a0 = c - u.x, a1 = c - u.y;
            if ((uint32_t)(2 * k) < ndev) key = a0 > key ? a0 : key;
            if ((uint32_t)(2 * k + 1) < ndev) key = a1 > key ? a1 : key;
        }
    } else {
        const unsigned long long* row = used + (size_t)node * dev_stride;
        for (uint32_t d = 0; d < ndev && d < dev_stride; ++d) {
            const unsigned long long a = c - row[d];
            key = a > key ? a : key;
        }
    }
    return key;
}

}  // namespace gs
