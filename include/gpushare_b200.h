/*
 * gpushare_b200.h -- C ABI of the B200-native gpushare scheduling hot path.
 *
 * This is the seam a maintainer of bnulwh/gpushare-scheduler-extender binds with cgo in place of the
 * in-memory `*cache.SchedulerCache` (pkg/cache/cache.go:14-28): the cluster lives on the GPU as a
 * struct-of-arrays snapshot, and the three hot loops of the reference run as CUDA kernels:
 *
 *   reference                                                     | entry point here
 *   --------------------------------------------------------------+----------------------------
 *   NewNodeInfo                  pkg/cache/nodeinfo.go:29-45       | gs_node_upsert
 *   BuildCache + DeviceInfo.GetUsedGPUMemory                       | gs_pods_load + gs_rebuild (K1)
 *                                pkg/cache/cache.go:49-74,         |
 *                                pkg/cache/deviceinfo.go:41-54     |
 *   AddOrUpdatePod / RemovePod   pkg/cache/cache.go:89-127,        | gs_pods_apply (K2)
 *     addPod / removePod         pkg/cache/deviceinfo.go:56-80     |
 *   Predicate.Handler loop over NodeInfo.Assume                    | gs_filter_batch (K3)
 *                                pkg/scheduler/predicate.go:15-39, |
 *                                pkg/cache/nodeinfo.go:113-137     |
 *   NodeInfo.allocateGPUID + the commit in Allocate                | gs_bind_batch (K4),
 *                                pkg/cache/nodeinfo.go:139-252     | gs_bind_rollback
 *   Inspect buildNode            pkg/scheduler/inspect.go:33-71    | gs_inspect
 *
 * Strings never cross this boundary: the host side keeps name->node_idx and (UID,node,dev)->slot
 * tables, parses annotations once at ingestion and renders the reference's reason strings.
 * Every integer is the reference's 64-bit Go int/uint; `uint` wrap-around is reproduced.
 *
 * Conventions: every function returns GS_OK (0) or a negative GS_ERR_* code and never throws or
 * aborts; gs_last_error() gives the message for the calling thread.  Host buffers belong to the
 * caller and are only touched during the call (cgo's pointer rule).  Each call makes the snapshot's
 * device current for the calling thread.
 *
 * Threads: any thread may call any function.  The Filter entry points (gs_filter_batch*,
 * gs_node_count) share the snapshot, as the reference's Filters share a node under RLock
 * (pkg/cache/nodeinfo.go:116): host-buffer Filters of different threads run side by side, each in
 * its own stream and work buffers.  Everything that changes the cluster state or reads it back
 * (upsert, load, rebuild, events, binds, rollback, inspect, read_*) is exclusive, like Allocate and
 * the pod events under the node's write lock (nodeinfo.go:72,90,141).  Different snapshots (one per
 * GPU) are independent.  Work an entry point leaves running on a CUDA stream (the *_device and
 * *_async calls) is ordered against later calls of the snapshot on the GPU, whatever stream they use.
 *
 * There is no CPU fallback: without a CUDA device gs_snapshot_create fails with GS_ERR_CUDA.
 */
#ifndef GPUSHARE_B200_H
#define GPUSHARE_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GS_ABI_VERSION 2

enum {
    GS_OK = 0,
    GS_ERR_INVALID = -1,  /* bad argument                                   */
    GS_ERR_CUDA = -2,     /* CUDA runtime error (message in gs_last_error)  */
    GS_ERR_CAPACITY = -3, /* index beyond what the snapshot was created for */
    GS_ERR_NOMEM = -4
};

/* Outcome classes of GetNodeInfo + IsGPUSharingNode (pkg/scheduler/gpushare-predicate.go:18-25). */
enum {
    GS_NODE_OK = 0,        /* in the node lister and gpu-mem capacity > 0                          */
    GS_NODE_NOT_FOUND = 1, /* `node "<name>" not found`                                            */
    GS_NODE_NOT_SHARE = 2  /* `The node <name> is not for GPU share, need skip`                    */
};

/* Most devices a node may have (len(NodeInfo.devs), pkg/cache/nodeinfo.go:32-35). */
#define GS_MAX_DEVS 64
/* Rows one filter / bind launch handles; larger batches are split internally. */
#define GS_MAX_BATCH 12288

/* One node as NewNodeInfo freezes it (pkg/cache/nodeinfo.go:29-45). */
typedef struct gs_node {
    uint64_t dev_capacity; /* uint(total / count), identical for every device of the node */
    uint32_t node_idx;
    uint32_t dev_count; /* gpu-count (<= 0 in the reference -> 0 here)                          */
    uint8_t node_class; /* GS_NODE_*                                                            */
    uint8_t pad_[7];
} gs_node;

#define GS_POD_VALID 1u   /* the slot holds an entry of some DeviceInfo.podMap                      */
#define GS_POD_COUNTED 2u /* phase is neither Succeeded nor Failed (pkg/cache/deviceinfo.go:46-49)  */

/* One podMap entry: pod -> (node, gpu-id, mem).  `mem` is GetGPUMemoryFromPodAnnotation
 * (pkg/utils/pod.go:94-113), `dev` GetGPUIDFromAnnotation (pod.go:45-60) clamped to int16. */
typedef struct gs_pod_record {
    uint64_t mem;
    uint32_t node_idx;
    int16_t dev;
    uint8_t flags; /* GS_POD_* */
    uint8_t pad_;
} gs_pod_record;

enum { GS_EV_ADD = 0, GS_EV_REMOVE = 1 };

/* One cache mutation.  ADD on a live slot replaces it (deviceinfo.go:63: same UID, same device);
 * REMOVE of a dead slot is a no-op.  Events of one batch apply in array order. */
typedef struct gs_pod_event {
    uint64_t mem;
    uint32_t slot;
    uint32_t node_idx;
    int16_t dev;
    uint8_t op;      /* GS_EV_* */
    uint8_t counted; /* 0/1, see GS_POD_COUNTED */
    uint32_t pad_;
} gs_pod_event;

/* One Bind.  `req` is uint(GetGPUMemoryFromPodResource) (nodeinfo.go:217); `commit_mem` is what the
 * committed pod's annotation will read back as (max(0, int(req)), pod.go:98-101); `slot` is a free
 * slot that receives the new podMap entry when a device is found. */
typedef struct gs_bind_req {
    uint64_t req;
    uint64_t commit_mem;
    uint32_t node_idx;
    uint32_t slot;
    uint8_t counted;
    uint8_t pad_[7];
} gs_bind_req;

#define GS_BIND_NO_DEVICE (-1)      /* allocateGPUID found nothing: "can't place the pod"        */
#define GS_BIND_NODE_NOT_FOUND (-2) /* GetNodeInfo failed for node_idx                           */
#define GS_BIND_SLOT_BUSY (-3)      /* `slot` holds an entry of another node: nothing was changed */

typedef struct gs_snapshot gs_snapshot;

/* Per-kernel device-time counters (CUDA events on the launching stream), for bench.py. */
enum {
    GS_K_REBUILD = 0,
    GS_K_APPLY = 1,
    GS_K_SORT = 2,
    GS_K_NODE_RANK = 3,
    GS_K_EXPAND = 4,
    GS_K_BIND_SORT = 5,
    GS_K_BIND_SCAN = 6,
    GS_K_COMPACT = 7,
    GS_K_SUBSET = 8,
    GS_K_COUNT_ = 9
};

const char* gs_last_error(void);
int gs_abi_version(void);
int gs_device_count(int* out_count);

/* Replaces NewSchedulerCache (pkg/cache/cache.go:30-38).  max_devs is the row stride of the
 * node x device arrays (1..GS_MAX_DEVS). */
int gs_snapshot_create(int device, uint32_t max_nodes, uint32_t max_devs, uint32_t max_pods,
                       gs_snapshot** out);
int gs_snapshot_destroy(gs_snapshot* s);

/* Replaces NewNodeInfo / the replacement branch of GetNodeInfo (cache.go:139-157).  Upserting a
 * node does NOT touch its used[] row; the caller removes the node's pods when the reference would
 * drop the podMaps. */
int gs_node_upsert(gs_snapshot* s, const gs_node* nodes, uint32_t n);
/* Nodes covered by filter bitmaps: 1 + the largest node_idx ever upserted. */
int gs_node_count(gs_snapshot* s, uint32_t* out_nodes);

/* Replaces the pod loop of BuildCache (cache.go:56-69): writes records into slots
 * [first_slot, first_slot + n).  Records whose (node, dev) is not a device of the snapshot are
 * stored invalid, as nodeinfo.go:98-108 would not add them.  Does not change used[]; follow with
 * gs_rebuild. */
int gs_pods_load(gs_snapshot* s, const gs_pod_record* recs, uint32_t first_slot, uint32_t n);
/* K1: used[node][dev] = sum of mem over valid, counted records (deviceinfo.go:41-54).  Reads every
 * slot of the pod table (max_pods; slots never written are invalid), whatever wrote them. */
int gs_rebuild(gs_snapshot* s);
/* K2: AddOrUpdatePod / RemovePod deltas (cache.go:89-127).  Returns when the events are applied. */
int gs_pods_apply(gs_snapshot* s, const gs_pod_event* events, uint32_t n);
/* The same, returning as soon as the batch is queued on the GPU (the events are copied before the call
 * returns); every later call of the snapshot sees the batch applied.  This is how the informer's work
 * queue (pkg/gpushare/controller.go:166-205) feeds the snapshot without stalling the Filter path.  A CUDA
 * error of the batch surfaces in a later call.  gs_snapshot_sync waits for everything queued so far. */
int gs_pods_apply_async(gs_snapshot* s, const gs_pod_event* events, uint32_t n);
int gs_snapshot_sync(gs_snapshot* s);

/* K3: the Predicate.Handler loop for a batch of pending pods against every node.
 * req[p] = uint(GetGPUMemoryFromPodResource(pod p)).  out_bitmap is [n_pods][row_stride_words]
 * 32-bit words, bit (n & 31) of word (n >> 5) of row p set iff the predicate passes for node n;
 * row_stride_words >= ceil(node_count / 32); padding words up to the next multiple of 8 (whole
 * 32-byte sectors), as far as they lie inside the row stride, may be written as zero.  A zero bit on a
 * GS_NODE_OK node means "Insufficient GPU Memory in one device".  Fastest when row_stride_words is a
 * multiple of 8 and out_bitmap came from gs_host_alloc. */
int gs_filter_batch(gs_snapshot* s, const uint64_t* req, uint32_t n_pods, uint32_t* out_bitmap,
                    uint32_t row_stride_words);
/* Same, with req and out_bitmap in device memory of the snapshot's GPU; asynchronous on
 * `cuda_stream` (a cudaStream_t; NULL = the snapshot's own stream -- to name CUDA's default stream pass
 * cudaStreamLegacy or cudaStreamPerThread).  The device-buffer calls of one snapshot share one set of
 * work buffers: the library orders them one after the other on the GPU, whatever streams they use. */
int gs_filter_batch_device(gs_snapshot* s, const uint64_t* d_req, uint32_t n_pods,
                           uint32_t* d_out_bitmap, uint32_t row_stride_words, void* cuda_stream);
/* K3 for the candidate list of one Filter request (`for _, nodeName := range *args.NodeNames`,
 * pkg/scheduler/predicate.go:21): bit (c & 31) of word (c >> 5) of row p is the predicate of pod p on node
 * node_subset[c].  An index that names no node of the snapshot, or a node that is not GS_NODE_OK, gives 0.
 * row_stride_words >= ceil(n_candidates / 32); padding words are written as zero.  Only
 * n_pods * ceil(n_candidates / 32) words cross PCIe, not a row of the whole cluster per pod. */
int gs_filter_batch_subset(gs_snapshot* s, const uint64_t* req, uint32_t n_pods, const uint32_t* node_subset,
                           uint32_t n_candidates, uint32_t* out_bitmap, uint32_t row_stride_words);
/* K3 in compact form.  The predicate is monotone in the request, so the [n_pods][nodes] answer is carried by
 * one number per node and one per pod:
 *     pass(p, n)  <=>  out_rank[(p / GS_MAX_BATCH) * rank_stride + n]  >=  out_cutoff[p]
 * exactly, for every 64-bit request (rank[n] = how many requests of the chunk the node can hold, 0 for a node
 * that can never pass; cutoff[p] = 1 + how many requests of the chunk are smaller than pod p's).  Ranks are
 * relative to a chunk of GS_MAX_BATCH pods: out_rank is [ceil(n_pods / GS_MAX_BATCH)][rank_stride] with
 * rank_stride >= node count (padding written as zero), out_cutoff is [n_pods].  2 bytes per node and per pod
 * cross PCIe instead of one bit per (pod, node); the caller expands what it needs.  When req, out_rank and
 * out_cutoff are page-locked (gs_host_alloc, or cudaHostAlloc / cudaHostRegister by the caller) and rank_stride
 * equals the node count, the kernels read and write them in place and no copy is set up; the call still returns
 * with the results in the caller's memory. */
int gs_filter_batch_compact(gs_snapshot* s, const uint64_t* req, uint32_t n_pods, uint16_t* out_rank,
                            uint32_t rank_stride, uint16_t* out_cutoff);
/* Same with device buffers, asynchronous on `cuda_stream` (padding of d_out_rank is left untouched). */
int gs_filter_batch_compact_device(gs_snapshot* s, const uint64_t* d_req, uint32_t n_pods, uint16_t* d_out_rank,
                                   uint32_t rank_stride, uint16_t* d_out_cutoff, void* cuda_stream);
/* The compact Filter of one node shard, gathered by the kernel that computes it: the shard's ranks go to
 * d_rank_dsts[k][dst_offset .. dst_offset + slice_count) for every k < n_dsts (<= 16) -- the nodes' ranks first, zeros up
 * to slice_count.  The destinations are device addresses valid in this process: this GPU's copy of the gathered
 * [shards][slice_count] array and the peers' copies mapped over NVLink (e.g. the buffer_ptrs of a torch symmetric-memory
 * allocation); the kernel reaches them with plain stores.  One chunk only (n_pods <= GS_MAX_BATCH).  The caller orders
 * the GPUs against each other afterwards (a barrier over the same peers) before any of them reads the gathered array.
 * SURVEY.md 8e's "NVLink gather of ... binpack winners" in its compact form; no counterpart in the reference. */
int gs_filter_batch_compact_device_gather(gs_snapshot* s, const uint64_t* d_req, uint32_t n_pods, uint16_t* const* d_rank_dsts,
                                          uint32_t n_dsts, uint32_t dst_offset, uint32_t slice_count, uint16_t* d_out_cutoff,
                                          void* cuda_stream);
/* Profiling aid: launches only the K3 stages named in stage_mask (GS_STAGE_*), in order, on the
 * same buffers gs_filter_batch_device uses (n_pods <= GS_MAX_BATCH).  Later stages read what earlier
 * stages left in the snapshot's work buffers, so run the full set once first.  bench.py times each
 * stage with CUDA events around a loop of such calls. */
#define GS_STAGE_SORT 1u
#define GS_STAGE_NODE_RANK 2u
#define GS_STAGE_EXPAND 4u
int gs_filter_stages_device(gs_snapshot* s, uint32_t stage_mask, const uint64_t* d_req, uint32_t n_pods,
                            uint32_t* d_out_bitmap, uint32_t row_stride_words, void* cuda_stream);
/* Per-node class bytes (GS_NODE_*) for rendering FailedNodes. */
int gs_node_classes(gs_snapshot* s, uint8_t* out_class, uint32_t n);

/* K4: allocateGPUID + commit for a queue of binds, applied in array order per node
 * (nodeinfo.go:139-252).  out_dev[i] = chosen device, GS_BIND_NO_DEVICE or GS_BIND_NODE_NOT_FOUND.
 * A found bind commits used[node][dev] += commit_mem (when counted) and fills `slot`.
 * Slots: a slot may be named by one bind of the call only (GS_ERR_INVALID otherwise, nothing is run);
 * a bind whose slot holds an entry of the SAME node replaces it (the same UID bound again,
 * deviceinfo.go:63); a slot that holds an entry of another node is refused: out_dev[i] =
 * GS_BIND_SLOT_BUSY, nothing changes (remove the entry first). */
int gs_bind_batch(gs_snapshot* s, const gs_bind_req* binds, uint32_t n, int32_t* out_dev);
/* Undo commits whose apiserver Update/Bind failed (nodeinfo.go:151-168,185-188: no commit). */
int gs_bind_rollback(gs_snapshot* s, const uint32_t* slots, uint32_t n);

/* Inspect (inspect.go:33-71): per-device capacity and used memory of one node. */
int gs_inspect(gs_snapshot* s, uint32_t node_idx, uint64_t* out_capacity, uint64_t* out_used,
               uint32_t cap_devs, uint32_t* out_dev_count, uint8_t* out_class);
/* Bulk read-back of used[first_node .. first_node+n)[max_devs] (row-major) and of pod records. */
int gs_read_used(gs_snapshot* s, uint32_t first_node, uint32_t n, uint64_t* out_used);
int gs_read_pods(gs_snapshot* s, uint32_t first_slot, uint32_t n, gs_pod_record* out);

/* Page-locked host memory for bitmaps / requests / ranks: copies run at full PCIe rate, and the compact Filter works
 * on such buffers in place.  Blocks from here are recognised without a driver call. */
int gs_host_alloc(size_t bytes, void** out);
int gs_host_free(void* p);

/* Options.  GS_OPT_PIPELINE_FILTERS (default 0): consecutive gs_filter_batch_device calls on one stream may
 * overlap on the GPU.  The calls rotate through separate work buffers; the second call's preparation kernel
 * runs beside the first call's bitmap kernel, and its bitmap kernel takes over the SM slots the first one frees
 * instead of waiting for it to drain (the kernels wait for each other through counters in device memory, not
 * through stream order).  Contract when set: the request array of a call must already be complete when the
 * PREVIOUS filter call was enqueued (do not produce it with a kernel enqueued in between; a cudaMemcpyAsync in
 * between is fine, it restores plain stream order), and a call whose output overlaps the previous call's output
 * is not overlapped.  Every other entry point ends a chain of overlapping filters.  Results are identical either
 * way; what the stream guarantees after a call (the bitmap is complete when the stream reaches that point) does
 * not change.  No counterpart in the reference. */
enum { GS_OPT_PIPELINE_FILTERS = 1 };
int gs_set_option(gs_snapshot* s, int option, int value);

/* Device-time profiling of the library's own kernels (off by default). */
int gs_profile_enable(gs_snapshot* s, int on);
int gs_profile_read(gs_snapshot* s, int kernel_id, uint64_t* out_launches, double* out_total_ms);
int gs_profile_reset(gs_snapshot* s);
/* Total kernels this snapshot has launched since creation (always counted). */
int gs_launch_count(gs_snapshot* s, uint64_t* out_launches);
/* SMs the kernels size their grids for: those of the CUDA context that was current when the snapshot was
 * created (fewer than the device has inside a green context or an MPS partition). */
int gs_sm_count(gs_snapshot* s, int* out_sms);
/* Diagnostics.  With overlapping Filter batches a CTA of the expansion kernel takes its batch's preparation kernel to be
 * complete from the order of the launches and only confirms it by an epoch it reads while its first inputs are on their
 * way; a CTA that finds the epoch not reached waits for it and asks again (the results are right either way).  This is
 * how many CTAs did so since the snapshot was created: 0 unless the launch order is not what the library produces.
 * Waits for the GPU. */
int gs_debug_early_starts(gs_snapshot* s, uint64_t* out_count);
/* Device pointers of the snapshot arrays, for zero-copy use from the same process (tests, NCCL). */
int gs_device_ptrs(gs_snapshot* s, void** d_used, void** d_capacity, void** d_nodeinfo, void** d_pods);
/* Tell the snapshot that work queued on `cuda_stream` by the caller (a kernel of its own, gs_pods_apply_static)
 * writes the arrays obtained from gs_device_ptrs: every later call of the snapshot is ordered behind that work,
 * and a chain of overlapping filters (GS_OPT_PIPELINE_FILTERS) ends here. */
int gs_snapshot_external_write(gs_snapshot* s, void* cuda_stream);

/* Pod events for callers that must not allocate.  (No counterpart in the reference: this exists only
 * because some hosts of this library cannot call cudaMalloc.)
 *
 * gs_pods_apply copies its events to the GPU and may grow a staging buffer, so it can call cudaMalloc /
 * cudaMallocHost.  A host that has given all GPU memory to one owner at start-up and forbids later
 * allocation -- or that sits on a path where the device-wide synchronisation hidden in cudaMalloc and
 * cudaFree is unacceptable -- uses this entry point instead: it calls no allocating routine at all.
 * The caller passes device pointers it already owns for everything:
 *   d_events   [n_events]            the batch, same meaning and order rules as gs_pods_apply
 *   d_pods     [n_slots]             the pod table, updated in place
 *   d_nodeinfo [n_nodes]             one word per node: dev_count | (GS_NODE_* class << 16)
 *   d_used     [n_nodes][dev_stride] used memory per device, updated in place
 * (the arrays of a snapshot, from gs_device_ptrs, qualify: follow the call with gs_snapshot_external_write
 * on the same stream so that the snapshot's own calls are ordered behind it).  An event takes effect only if
 * it is a GS_EV_REMOVE, or a GS_EV_ADD onto a device index below both the node's dev_count and dev_stride;
 * other op codes are skipped.  The only scratch -- one int32 per pod slot
 * -- is a fixed array of GS_STATIC_MAX_SLOTS entries (4 MiB) inside the library; every GPU that loads
 * the library holds it until the process exits.  n_slots above GS_STATIC_MAX_SLOTS is refused with
 * GS_ERR_CAPACITY.
 *
 * The first call on a device looks up the array's device address (cudaGetSymbolAddress, once per
 * device); make that first call before any CUDA stream capture begins.  The work is queued on
 * `cuda_stream` (NULL = the default stream) and the call returns without waiting for it.  All calls on
 * one GPU share the one array, which keeps what the last call left in it: the entry point resets the
 * part it is about to use at the start of every call, and two calls whose kernels overlap on the GPU
 * (different streams, different threads) race on it -- order them on one stream or with events. */
#define GS_STATIC_MAX_SLOTS (1u << 20)
int gs_pods_apply_static(int device, const gs_pod_event* d_events, uint32_t n_events, gs_pod_record* d_pods,
                         uint32_t n_slots, const uint32_t* d_nodeinfo, uint32_t n_nodes, uint32_t dev_stride,
                         uint64_t* d_used, void* cuda_stream);

#ifdef __cplusplus
}
#endif
#endif
