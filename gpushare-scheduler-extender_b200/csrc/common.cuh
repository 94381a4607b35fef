// Shared definitions of the snapshot object and launch helpers (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>
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
constexpr int kTileRows = 64;            // sorted pod rows per tile
constexpr int kTileWords = 128;          // 32-bit words (4096 nodes) per tile
constexpr int kExpandThreads = 256;

struct ProfSlot {
    uint64_t launches = 0;
    double total_ms = 0.0;
};

}  // namespace gs

struct gs_snapshot {
    int device = 0;
    uint32_t max_nodes = 0, max_devs = 0, max_pods = 0;
    uint32_t n_nodes = 0;   // 1 + largest node_idx upserted
    uint32_t pods_hw = 0;   // 1 + largest slot ever written
    int sm_count = 148;
    cudaStream_t stream = nullptr;

    // cluster state (HBM)
    unsigned long long* d_used = nullptr;  // [max_nodes][max_devs]
    unsigned long long* d_cap = nullptr;   // [max_nodes]
    uint32_t* d_nodeinfo = nullptr;        // [max_nodes]
    gs_pod_record* d_pods = nullptr;       // [max_pods]

    // work buffers (HBM scratch owned by the snapshot)
    int32_t* d_last_ev = nullptr;             // [max_pods], -1 when idle (K2)
    unsigned long long* d_sorted_keys = nullptr;  // [GS_MAX_BATCH] (K3/K4 sort output)
    uint16_t* d_perm = nullptr;               // [GS_MAX_BATCH] sorted position -> input index
    uint16_t* d_row_bucket = nullptr;         // [GS_MAX_BATCH] (K3)
    uint32_t* d_pod_hist = nullptr;           // [2 sets][512] bucket counts (K3, self re-arming)
    unsigned long long max_capacity = 0;      // largest dev_capacity of an OK node ever upserted (bucket scale)
    uint32_t* d_rank_planes = nullptr;        // [rank_words / kTileWords][kRankPlanes][kTileWords] (K3)
    unsigned long long* d_node_key = nullptr; // [rank_words * 32] (K3)
    uint32_t* d_node_key32 = nullptr;         // [rank_words * 32] (K3)
    uint32_t* d_sched = nullptr;              // [2] tile scheduler of k_expand (self re-arming)
    // The K3 work buffers above exist twice (set 0 / set 1, second copy right behind the first) and filter
    // launches alternate between them, so that with GS_OPT_PIPELINE_FILTERS the prep kernel of one batch can run
    // under the tail of the previous batch's expand kernel.
    uint32_t fb_set = 0;
    uint32_t* d_epochs = nullptr;     // per set s: [3 * s] finished k_prep blocks, [3 * s + 1] k_prep epoch, [3 * s + 2] k_expand epoch
    uint32_t n_prep[2] = {0, 0};      // k_prep / k_expand launches issued per set (what the epochs will reach)
    uint32_t n_expand[2] = {0, 0};
    bool pipeline = false;            // GS_OPT_PIPELINE_FILTERS
    bool pipeline_dev = false;        // development builds: expand-only launches overlap as well
    bool prev_filter = false;         // the last operation enqueued for this snapshot was a filter ...
    cudaStream_t prev_stream = nullptr;  // ... on this stream ...
    const char* prev_out_lo = nullptr;   // ... writing [prev_out_lo, prev_out_hi)
    const char* prev_out_hi = nullptr;
    uint32_t rank_words = 0;
    void* d_stage = nullptr;                  // generic device staging (events, binds, reqs, out)
    size_t d_stage_bytes = 0;
    void* d_bitmap = nullptr;                 // device bitmap for the host-buffer filter variant
    size_t d_bitmap_bytes = 0;
    void* h_stage = nullptr;                  // pinned host staging
    size_t h_stage_bytes = 0;

    // profiling
    bool profiling = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    gs::ProfSlot prof[GS_K_COUNT_];
    uint64_t launches = 0;

    std::mutex mu;
};
