// Shared definitions of the snapshot object and launch helpers (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <condition_variable>
#include <mutex>
#include <shared_mutex>
#include <string>

#include "gpushare_b200.h"
#include "kernels.h"

namespace gs {

// nodeinfo word: low 16 bits = device count, bits 16..23 = GS_NODE_* class.
__host__ __device__ inline uint32_t pack_nodeinfo(uint32_t ndev, uint32_t cls) { return (ndev & 0xffffu) | (cls << 16); }
__host__ __device__ inline uint32_t nodeinfo_ndev(uint32_t w) { return w & 0xffffu; }
__host__ __device__ inline uint32_t nodeinfo_class(uint32_t w) { return (w >> 16) & 0xffu; }

constexpr uint32_t kUnsetNodeinfo = (uint32_t)GS_NODE_NOT_FOUND << 16;

// Filter geometry.
constexpr int kRankPlanes = 10;          // bit planes per 32-node word of the node bucket ids (0..513)
constexpr int kTileRows = 64;            // sorted pod rows per tile (32: 18.3 us alone, 128: 20.5 us; profiles/NOTES.md 27)
constexpr int kDensePlanes = 7;          // bit planes of a pass count 0..kTileRows
constexpr int kTileWords = 128;          // 32-bit words (4096 nodes) per tile
constexpr int kExpandThreads = 256;         // (128-thread CTAs, 4-6 per SM: 20-22 us alone; profiles/NOTES.md 28)


// Reader/writer lock of the snapshot with Go's sync.RWMutex rule: once a writer waits, new readers queue behind
// it, so a stream of Filters cannot starve the pod events (pthread's default prefers readers).
class RwLock {
public:
    void lock_shared() {
        std::unique_lock<std::mutex> l(m_);
        cv_.wait(l, [&] { return !writer_ && writers_waiting_ == 0; });
        ++readers_;
    }
    void unlock_shared() {
        std::lock_guard<std::mutex> l(m_);
        if (--readers_ == 0) cv_.notify_all();
    }
    void lock() {
        std::unique_lock<std::mutex> l(m_);
        ++writers_waiting_;
        cv_.wait(l, [&] { return !writer_ && readers_ == 0; });
        --writers_waiting_;
        writer_ = true;
    }
    void unlock() {
        std::lock_guard<std::mutex> l(m_);
        writer_ = false;
        cv_.notify_all();
    }

private:
    std::mutex m_;
    std::condition_variable cv_;
    int readers_ = 0, writers_waiting_ = 0;
    bool writer_ = false;
};

struct ProfSlot {
    uint64_t launches = 0;
    double total_ms = 0.0;
};

// Everything one in-flight Filter needs besides the cluster state: a stream, staging, and the K3 work
// buffers.  A snapshot owns a few of these so that Filter calls from different host threads run side by
// side, as the reference's Filters do under the node read locks (pkg/cache/nodeinfo.go:116).
struct FilterCtx {
    std::mutex mu;                       // one call at a time uses a context
    cudaStream_t stream = nullptr;       // the context's own stream (host-buffer calls)
    cudaEvent_t done = nullptr;          // recorded, when a writer needs it, behind the context's outstanding work
    cudaStream_t last_stream = nullptr;  // where the context's last work was enqueued
    bool outstanding = false;            // that work may still be running (asynchronous device-buffer calls)
    uint64_t seen_write_seq = 0;         // the snapshot's write_seq this context's streams are ordered after

    // K3 work buffers; all of them exist twice (set 0 / set 1, second copy right behind the first): launches
    // alternate between the sets so that with GS_OPT_PIPELINE_FILTERS the prep kernel of one batch can run
    // under the tail of the previous batch's expand kernel.
    unsigned long long* d_grouped_req = nullptr;  // [2][GS_MAX_BATCH]
    uint16_t* d_perm = nullptr;                   // [2][GS_MAX_BATCH] grouped position -> input index
    uint16_t* d_row_bucket = nullptr;             // [2][GS_MAX_BATCH]
    uint32_t* d_pod_hist = nullptr;               // [2][512] bucket counts (self re-arming)
    uint32_t* d_rank_planes = nullptr;            // [2][rank_words / kTileWords][kRankPlanes][kTileWords]
    unsigned long long* d_node_key = nullptr;     // [2][rank_words * 32]
    uint32_t* d_node_key32 = nullptr;             // [2][rank_words * 32]
    uint32_t* d_sched = nullptr;                  // [2][2] tile scheduler of k_expand (self re-arming)
    // per set s: [3 * s] finished k_prep blocks (all launches), [3 * s + 1] k_prep epoch, [3 * s + 2] k_expand epoch;
    // [6] CTAs of k_expand that started before their k_prep's blocks had all finished (diagnostic)
    uint32_t* d_epochs = nullptr;
    uint32_t fb_set = 0;
    uint32_t n_prep_blocks[2] = {0, 0};  // what the set's count of finished k_prep blocks will reach
    uint32_t n_prep[2] = {0, 0};      // what the set's k_prep / k_expand epochs will reach when everything issued has run
    uint32_t n_expand[2] = {0, 0};
    bool prev_filter = false;            // the last operation enqueued for this context was a full filter ...
    cudaStream_t prev_stream = nullptr;  // ... on this stream ...
    const char* prev_out_lo = nullptr;   // ... writing [prev_out_lo, prev_out_hi)
    const char* prev_out_hi = nullptr;

    // compact filter (rank / cutoff form): requests sorted slice by slice, fully sorted requests, duplicate counters
    unsigned long long* d_cslices = nullptr;    // [GS_MAX_BATCH rounded up to whole slices of 1024]
    unsigned long long* d_csorted = nullptr;    // [GS_MAX_BATCH]
    uint32_t* d_cdup = nullptr;                 // [GS_MAX_BATCH] zero between calls (self re-arming)
    unsigned long long* d_creq = nullptr;       // [GS_MAX_BATCH] device copy of requests that live in host memory
    uint16_t* d_ccut = nullptr;                 // [GS_MAX_BATCH] cutoffs on their way to host memory

    void* d_stage = nullptr;   // device staging of host-buffer calls (requests, candidate lists, compact results)
    size_t d_stage_bytes = 0;
    void* d_bitmap = nullptr;  // device bitmap of host-buffer calls
    size_t d_bitmap_bytes = 0;
    void* h_stage = nullptr;   // pinned host staging for pageable caller buffers
    size_t h_stage_bytes = 0;
};

constexpr int kFilterCtxs = 4;
constexpr int kApplyRing = 2;  // pinned event staging slots of gs_pods_apply_async

}  // namespace gs

struct gs_snapshot {
    int device = 0;
    uint32_t max_nodes = 0, max_devs = 0, max_pods = 0;
    uint32_t n_nodes = 0;   // 1 + largest node_idx upserted
    int sm_count = 148;     // SMs of the context the snapshot was created in
    cudaStream_t stream = nullptr;  // the writers' stream (events, binds, rebuild, upserts)

    // cluster state (HBM)
    unsigned long long* d_used = nullptr;  // [max_nodes][max_devs]
    unsigned long long* d_cap = nullptr;   // [max_nodes]
    uint32_t* d_nodeinfo = nullptr;        // [max_nodes]
    gs_pod_record* d_pods = nullptr;       // [max_pods]

    // work buffers of the writers (HBM scratch owned by the snapshot)
    int32_t* d_last_ev = nullptr;                // [max_pods], -1 when idle (K2)
    unsigned long long* d_bind_keys = nullptr;   // [GS_MAX_BATCH] (K4 sort output)
    uint16_t* d_bind_perm = nullptr;             // [GS_MAX_BATCH] sorted position -> input index
    unsigned long long max_capacity = 0;         // largest dev_capacity of an OK node ever upserted (bucket scale)
    uint32_t rank_words = 0;
    void* d_stage = nullptr;                  // device staging (events, binds, results)
    size_t d_stage_bytes = 0;
    void* h_stage = nullptr;                  // pinned host staging
    size_t h_stage_bytes = 0;
    // gs_pods_apply_async: pinned + device staging slots, each with the event that says "slot is free again"
    void* h_ring[gs::kApplyRing] = {};
    void* d_ring[gs::kApplyRing] = {};
    size_t ring_bytes[gs::kApplyRing] = {};
    cudaEvent_t ring_free[gs::kApplyRing] = {};
    bool ring_used[gs::kApplyRing] = {};
    uint32_t ring_next = 0;

    // Readers (filters, read-backs) share the lock, writers take it exclusively: pkg/cache/nodeinfo.go:116,141.
    gs::RwLock state_mu;
    // Ordering on the device between the writers' stream and the filter contexts' streams.
    cudaEvent_t ev_write = nullptr;   // recorded on `stream` behind the last asynchronous write
    uint64_t write_seq = 0;           // bumped by every asynchronous write (under the exclusive lock)
    gs::FilterCtx ctx[gs::kFilterCtxs];
    std::atomic<uint32_t> ctx_rr{0};

    bool pipeline = false;            // GS_OPT_PIPELINE_FILTERS (context 0, device-buffer calls)
    bool pipeline_dev = false;        // development builds: expand-only launches overlap as well
    unsigned int* h_fault = nullptr;  // mapped pinned word: a kernel that gives up a bounded wait says where
    unsigned int* d_fault = nullptr;  // its device alias

    // profiling
    std::mutex prof_mu;
    bool profiling = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    gs::ProfSlot prof[GS_K_COUNT_];
    std::atomic<uint64_t> launches{0};
};
