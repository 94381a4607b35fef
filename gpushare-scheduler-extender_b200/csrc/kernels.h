// Launchers shared between the kernel translation units and the C-ABI layer.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "gpushare_b200.h"

namespace gs {

// message returned by gs_last_error() on the calling thread (api.cu)
void set_last_error(const char* msg);

// K3 (filter_kernels.cu)
struct FilterBuffers {              // work buffers of one filter launch sequence (HBM scratch)
    unsigned long long* grouped_req;  // [GS_MAX_BATCH] requests in bucket order
    uint16_t* perm;                   // [GS_MAX_BATCH] grouped position -> pod index
    uint16_t* row_bucket;             // [GS_MAX_BATCH] 1-based bucket of every grouped row
    uint32_t* pod_hist;               // [512] bucket counts of this buffer set (zero between launches)
    uint32_t* planes;                 // [words / kTileWords][kRankPlanes][kTileWords] node bucket ids, bit-sliced
    unsigned long long* node_key;     // [words * 32] max available memory per node
    uint32_t* node_key32;             // [words * 32] the same, saturated to 32 bits
    uint32_t* sched;                  // [2] tile scheduler of k_expand (self re-arming)
    uint32_t* prep_done;              // [1] finished k_prep blocks of all launches on this buffer set (never reset)
    uint32_t* prep_epoch;             // [1] completed k_prep launches on this buffer set
    uint32_t* expand_epoch;           // [1] completed k_expand launches on this buffer set
    uint32_t set_index;               // 0 / 1: which of the two buffer sets this is (development traces)
    unsigned int* fault;              // mapped host word: where a kernel gave up a bounded wait (may be null)
    uint32_t* early_starts;           // [1] CTAs of k_expand that found their batch's k_prep unfinished (expected: 0)
};
cudaError_t launch_sort_binds(int device, const gs_bind_req* d_binds, uint32_t n, unsigned long long* d_sorted,
                              uint16_t* d_perm, cudaStream_t st);
cudaError_t launch_prep(const unsigned long long* d_req, uint32_t n_pods, bool do_pods, bool do_nodes, uint32_t shift,
                        const FilterBuffers& fb, const unsigned long long* d_used, const unsigned long long* d_cap,
                        const uint32_t* d_nodeinfo, uint32_t n_nodes, uint32_t dev_stride, uint32_t n_words,
                        bool overlap_previous, uint32_t expand_epoch_wanted, uint32_t* blocks_total,
                        cudaStream_t st);
cudaError_t launch_expand(int sm_count, const FilterBuffers& fb, uint32_t n_pods, uint32_t words,
                          uint32_t row_stride, uint32_t* d_out, bool overlap_previous, uint32_t prep_epoch_wanted,
                          uint32_t prep_blocks_wanted, cudaStream_t st);

// K1 / K2 / K4 (state_kernels.cu)
cudaError_t launch_validate_records(gs_pod_record* d_pods, uint32_t first, uint32_t n, const uint32_t* d_nodeinfo,
                                    uint32_t n_nodes, uint32_t dev_stride, cudaStream_t st);
// zeroes used[0 .. n_node_rows * dev_stride) and adds up the table: two launches
cudaError_t launch_rebuild(const gs_pod_record* d_pods, uint32_t n_slots, uint32_t n_node_rows, uint32_t dev_stride,
                           unsigned long long* d_used, int sm_count, cudaStream_t st);
cudaError_t launch_apply_events(const gs_pod_event* d_events, uint32_t n, gs_pod_record* d_pods, uint32_t max_pods,
                                int32_t* d_last_ev, const uint32_t* d_nodeinfo, uint32_t n_nodes,
                                uint32_t dev_stride, unsigned long long* d_used, cudaStream_t st);
cudaError_t launch_bind_small(const gs_bind_req* d_binds, uint32_t n, gs_pod_record* d_pods, uint32_t max_pods,
                              const uint32_t* d_nodeinfo, uint32_t n_nodes, const unsigned long long* d_cap, uint32_t dev_stride,
                              unsigned long long* d_used, int32_t* d_out_dev, cudaStream_t st);
cudaError_t launch_bind_scan(const gs_bind_req* d_binds, uint32_t n, const unsigned long long* d_sorted_nodes,
                             const uint16_t* d_perm, gs_pod_record* d_pods, uint32_t max_pods,
                             const uint32_t* d_nodeinfo, uint32_t n_nodes, const unsigned long long* d_cap,
                             uint32_t dev_stride, unsigned long long* d_used, int32_t* d_out_dev, cudaStream_t st);
cudaError_t launch_fill_i32(int32_t* d, int32_t v, size_t n, cudaStream_t st);

// compact and candidate-list forms of K3 (compact_kernels.cu)
// Where k_node_rank writes the ranks: one destination (the caller's buffer) or several -- this GPU's slice of a
// gathered [ranks][count] array on every GPU of the box, the peers' copies reached through NVLink by plain stores
// (the all-gather of the sharded compact Filter, done by the kernel that produces the data).
constexpr int kMaxRankDsts = 16;
struct RankDsts {
    uint16_t* p[kMaxRankDsts];  // base of each destination array
    uint32_t n;                 // destinations in use
    uint32_t offset;            // first element written in each (this shard's slice)
    uint32_t count;             // elements written: the nodes, then zeros up to here
};
// rank[n] = #{p : req[p] <= key[n]} (0 for a node that can never pass), cutoff[p] = 1 + #{q : req[q] < req[p]}:
// pass(p, n) <=> rank[n] >= cutoff[p].  Three launches; returns how many were made through *n_launches.
cudaError_t launch_compact(const unsigned long long* d_req, uint32_t n_pods, unsigned long long* d_slices, unsigned long long* d_sorted,
                           uint32_t* d_dup, const unsigned long long* d_used, const unsigned long long* d_cap,
                           const uint32_t* d_nodeinfo, uint32_t n_nodes, uint32_t dev_stride, const RankDsts& dsts,
                           uint16_t* d_cutoff, int sm_count, cudaStream_t st, uint32_t* n_launches,
                           unsigned long long* d_req_copy = nullptr, uint16_t* d_cut_tmp = nullptr);
// bit c of row p = predicate(pod p, node subset[c]); one launch
cudaError_t launch_filter_subset(const unsigned long long* d_req, uint32_t n_pods, const uint32_t* d_subset, uint32_t n_cand,
                                 const unsigned long long* d_used, const unsigned long long* d_cap,
                                 const uint32_t* d_nodeinfo, uint32_t n_nodes, uint32_t dev_stride, uint32_t* d_out,
                                 uint32_t row_stride, cudaStream_t st);

}  // namespace gs
