// Launchers shared between the kernel translation units and the C-ABI layer.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "gpushare_b200.h"

namespace gs {

// K3 (filter_kernels.cu)
cudaError_t launch_sort_requests(int device, const unsigned long long* d_req, uint32_t n,
                                 unsigned long long* d_sorted, uint16_t* d_perm, cudaStream_t st);
cudaError_t launch_sort_binds(int device, const gs_bind_req* d_binds, uint32_t n, unsigned long long* d_sorted,
                              uint16_t* d_perm, cudaStream_t st);
cudaError_t launch_node_rank(const unsigned long long* d_used, const unsigned long long* d_cap,
                             const uint32_t* d_nodeinfo, uint32_t n_nodes, uint32_t dev_stride,
                             const unsigned long long* d_sorted, uint32_t n_pods, uint32_t* d_planes,
                             uint32_t n_words, cudaStream_t st);
cudaError_t launch_expand(int device, int sm_count, const uint32_t* d_planes, const uint16_t* d_perm,
                          uint32_t n_pods, uint32_t words, uint32_t row_stride, uint32_t* d_out, cudaStream_t st);

// K1 / K2 / K4 (state_kernels.cu)
cudaError_t launch_validate_records(gs_pod_record* d_pods, uint32_t first, uint32_t n, const uint32_t* d_nodeinfo,
                                    uint32_t n_nodes, cudaStream_t st);
cudaError_t launch_rebuild(const gs_pod_record* d_pods, uint32_t n_slots, uint32_t dev_stride,
                           unsigned long long* d_used, int sm_count, cudaStream_t st);
cudaError_t launch_apply_events(const gs_pod_event* d_events, uint32_t n, gs_pod_record* d_pods, uint32_t max_pods,
                                int32_t* d_last_ev, const uint32_t* d_nodeinfo, uint32_t n_nodes,
                                uint32_t dev_stride, unsigned long long* d_used, cudaStream_t st);
cudaError_t launch_bind_scan(const gs_bind_req* d_binds, uint32_t n, const unsigned long long* d_sorted_nodes,
                             const uint16_t* d_perm, gs_pod_record* d_pods, uint32_t max_pods,
                             const uint32_t* d_nodeinfo, uint32_t n_nodes, const unsigned long long* d_cap,
                             uint32_t dev_stride, unsigned long long* d_used, int32_t* d_out_dev, cudaStream_t st);
cudaError_t launch_fill_i32(int32_t* d, int32_t v, size_t n, cudaStream_t st);

}  // namespace gs
