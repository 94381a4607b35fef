// C-ABI layer (include/gpushare_b200.h) over the snapshot object and the kernel launchers.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <vector>

#include "common.cuh"
#include "kernels.h"

using u64 = unsigned long long;

static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}
static int cuda_fail(cudaError_t e, const char* what) {
    return fail(GS_ERR_CUDA, "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
}
#define CU(call)                                        \
    do {                                                \
        cudaError_t e_ = (call);                        \
        if (e_ != cudaSuccess) return cuda_fail(e_, #call); \
    } while (0)

namespace gs {
void set_last_error(const char* msg) { snprintf(g_err, sizeof g_err, "%s", msg); }
}  // namespace gs

namespace {

struct Guard {
    gs_snapshot* s;
    std::unique_lock<std::mutex> lk;
    // every entry point except the filters breaks a chain of back-to-back filters (see filter_chunk)
    explicit Guard(gs_snapshot* s_, bool is_filter = false) : s(s_), lk(s_->mu) {
        if (!is_filter) s->prev_filter = false;
    }
};

int use_device(gs_snapshot* s) {
    CU(cudaSetDevice(s->device));
    return GS_OK;
}

int ensure_dev_stage(gs_snapshot* s, size_t bytes) {
    if (bytes <= s->d_stage_bytes) return GS_OK;
    size_t want = std::max(bytes, s->d_stage_bytes * 2);
    if (s->d_stage) CU(cudaFree(s->d_stage));
    s->d_stage = nullptr;
    s->d_stage_bytes = 0;
    CU(cudaMalloc(&s->d_stage, want));
    s->d_stage_bytes = want;
    return GS_OK;
}
int ensure_bitmap(gs_snapshot* s, size_t bytes) {
    if (bytes <= s->d_bitmap_bytes) return GS_OK;
    if (s->d_bitmap) CU(cudaFree(s->d_bitmap));
    s->d_bitmap = nullptr;
    s->d_bitmap_bytes = 0;
    CU(cudaMalloc(&s->d_bitmap, bytes));
    s->d_bitmap_bytes = bytes;
    return GS_OK;
}
int ensure_host_stage(gs_snapshot* s, size_t bytes) {
    if (bytes <= s->h_stage_bytes) return GS_OK;
    size_t want = std::max(bytes, s->h_stage_bytes * 2);
    if (s->h_stage) CU(cudaFreeHost(s->h_stage));
    s->h_stage = nullptr;
    s->h_stage_bytes = 0;
    CU(cudaMallocHost(&s->h_stage, want));
    s->h_stage_bytes = want;
    return GS_OK;
}

bool is_pinned(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

// Host -> device copy of a caller buffer that may be pageable: pageable data is first copied into
// the snapshot's pinned staging so the DMA runs at full rate and asynchronously.
int h2d(gs_snapshot* s, void* dst, const void* src, size_t bytes) {
    if (bytes == 0) return GS_OK;
    if (is_pinned(src)) {
        CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, s->stream));
        return GS_OK;
    }
    int rc = ensure_host_stage(s, bytes);
    if (rc) return rc;
    CU(cudaStreamSynchronize(s->stream));  // staging may still be in flight from an earlier copy
    memcpy(s->h_stage, src, bytes);
    CU(cudaMemcpyAsync(dst, s->h_stage, bytes, cudaMemcpyHostToDevice, s->stream));
    return GS_OK;
}
// Device -> host into a caller buffer; ends synchronised.
int d2h(gs_snapshot* s, void* dst, const void* src, size_t bytes) {
    if (bytes == 0) return GS_OK;
    if (is_pinned(dst)) {
        CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, s->stream));
        CU(cudaStreamSynchronize(s->stream));
        return GS_OK;
    }
    int rc = ensure_host_stage(s, bytes);
    if (rc) return rc;
    CU(cudaMemcpyAsync(s->h_stage, src, bytes, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    memcpy(dst, s->h_stage, bytes);
    return GS_OK;
}

// Runs one launcher, optionally bracketed by events for the per-kernel counters.
template <class F>
int timed(gs_snapshot* s, int kid, cudaStream_t st, uint64_t n_launches, F&& f) {
    if (s->profiling) CU(cudaEventRecord(s->ev0, st));
    cudaError_t e = f();
    if (e != cudaSuccess) return cuda_fail(e, "kernel launch");
    s->launches += n_launches;
    if (s->profiling) {
        CU(cudaEventRecord(s->ev1, st));
        CU(cudaEventSynchronize(s->ev1));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
        s->prof[kid].launches += n_launches;
        s->prof[kid].total_ms += ms;
    }
    return GS_OK;
}

uint32_t bitmap_words(const gs_snapshot* s) { return (s->n_nodes + 31u) / 32u; }

// The three filter kernels for one chunk of <= GS_MAX_BATCH pods; everything on `st`.
int filter_chunk(gs_snapshot* s, const u64* d_req, uint32_t n_pods, uint32_t* d_out, uint32_t row_stride,
                 cudaStream_t st, uint32_t stages = 7u) {
    const uint32_t words = bitmap_words(s);
    if (n_pods == 0 || words == 0) return GS_OK;
    const uint32_t words_pad = (words + gs::kTileWords - 1) / gs::kTileWords * gs::kTileWords;
    if (words_pad > s->rank_words) return fail(GS_ERR_CAPACITY, "rank planes too small");
    const uint32_t set = s->fb_set;
    // full launches alternate; single-stage profiling launches stay on one set (development builds can let
    // expand-only launches alternate and overlap too: GS_EXPAND_TRACE + option value 2)
    const bool dev_overlap_expand = s->pipeline_dev && stages == GS_STAGE_EXPAND;
    if (stages == 7u || dev_overlap_expand) s->fb_set ^= 1u;
    gs::FilterBuffers fb;
    fb.grouped_req = s->d_sorted_keys + (size_t)set * GS_MAX_BATCH;
    fb.perm = s->d_perm + (size_t)set * GS_MAX_BATCH;
    fb.row_bucket = s->d_row_bucket + (size_t)set * GS_MAX_BATCH;
    fb.pod_hist = s->d_pod_hist + set * 512;
    fb.planes = s->d_rank_planes + (size_t)set * s->rank_words * gs::kRankPlanes;
    fb.node_key = s->d_node_key + (size_t)set * s->rank_words * 32;
    fb.node_key32 = s->d_node_key32 + (size_t)set * s->rank_words * 32;
    fb.sched = s->d_sched + set * 2;
    fb.set_index = set;
    fb.prep_done = s->d_epochs + 3 * set;
    fb.prep_epoch = s->d_epochs + 3 * set + 1;
    fb.expand_epoch = s->d_epochs + 3 * set + 2;
    // With GS_OPT_PIPELINE_FILTERS, a filter that directly follows another filter of this snapshot on the same
    // stream, writing somewhere else, does not wait for that filter's expand kernel to drain before its own prep
    // kernel starts (the two touch different work-buffer sets).  Anything else keeps plain stream order.
    const char* out_lo = reinterpret_cast<const char*>(d_out);
    const char* out_hi = out_lo + (size_t)n_pods * row_stride * sizeof(uint32_t);
    const bool overlap_ok = dev_overlap_expand ||
                            (s->pipeline && stages == 7u && s->prev_filter && s->prev_stream == st &&
                             (out_hi <= s->prev_out_lo || out_lo >= s->prev_out_hi));
    // bucket scale: requests up to the largest device capacity spread over all 512 buckets
    uint32_t bits = 0;
    while (bits < 64 && (s->max_capacity >> bits) != 0) ++bits;
    const uint32_t shift = bits > 9 ? bits - 9 : 0;
    int rc = GS_OK;
    const bool do_pods = (stages & GS_STAGE_SORT) != 0, do_nodes = (stages & GS_STAGE_NODE_RANK) != 0;
    if (do_pods || do_nodes)
        rc = timed(s, do_pods ? GS_K_SORT : GS_K_NODE_RANK, st, 1, [&] {
            cudaError_t e = gs::launch_prep(d_req, n_pods, do_pods, do_nodes, shift, fb, s->d_used, s->d_cap, s->d_nodeinfo,
                                            s->n_nodes, s->max_devs, words_pad, overlap_ok, s->n_expand[set], st);
            if (e == cudaSuccess) ++s->n_prep[set];
            return e;
        });
    if (rc) return rc;
    if (stages & GS_STAGE_EXPAND)
        rc = timed(s, GS_K_EXPAND, st, 1,
                   [&] {
                       cudaError_t e = gs::launch_expand(s->sm_count, fb, n_pods, words, row_stride, d_out, overlap_ok,
                                                         s->n_prep[set], st);
                       if (e == cudaSuccess) ++s->n_expand[set];
                       return e;
                   });
    s->prev_filter = rc == GS_OK && stages == 7u && !s->profiling;
    s->prev_stream = st;
    s->prev_out_lo = out_lo;
    s->prev_out_hi = out_hi;
    return rc;
}

int filter_device_locked(gs_snapshot* s, const u64* d_req, uint32_t n_pods, uint32_t* d_out, uint32_t row_stride,
                         cudaStream_t st) {
    if (row_stride < bitmap_words(s)) return fail(GS_ERR_INVALID, "row_stride_words %u < %u", row_stride, bitmap_words(s));
    for (uint32_t p0 = 0; p0 < n_pods; p0 += GS_MAX_BATCH) {
        const uint32_t n = std::min<uint32_t>(GS_MAX_BATCH, n_pods - p0);
        int rc = filter_chunk(s, d_req + p0, n, d_out + (size_t)p0 * row_stride, row_stride, st);
        if (rc) return rc;
    }
    return GS_OK;
}

}  // namespace

extern "C" {

const char* gs_last_error(void) { return g_err; }
int gs_abi_version(void) { return GS_ABI_VERSION; }

int gs_device_count(int* out_count) {
    if (!out_count) return fail(GS_ERR_INVALID, "null out_count");
    *out_count = 0;
    CU(cudaGetDeviceCount(out_count));
    return GS_OK;
}

int gs_snapshot_create(int device, uint32_t max_nodes, uint32_t max_devs, uint32_t max_pods, gs_snapshot** out) {
    if (!out) return fail(GS_ERR_INVALID, "null out");
    *out = nullptr;
    if (max_nodes == 0 || max_devs == 0 || max_devs > GS_MAX_DEVS)
        return fail(GS_ERR_INVALID, "max_nodes must be > 0 and max_devs in 1..%d", GS_MAX_DEVS);
    if ((uint64_t)max_nodes * max_devs >= (1ull << 32)) return fail(GS_ERR_INVALID, "max_nodes * max_devs must be < 2^32");
    if (max_pods == 0) max_pods = 1;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        cudaGetLastError();
        return fail(GS_ERR_CUDA, "no CUDA device: the gpushare B200 path has no CPU fallback (%s)",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    }
    if (device < 0 || device >= count) return fail(GS_ERR_INVALID, "device %d out of range (0..%d)", device, count - 1);
    CU(cudaSetDevice(device));
    gs_snapshot* s = new (std::nothrow) gs_snapshot();
    if (!s) return fail(GS_ERR_NOMEM, "out of host memory");
    s->device = device;
    s->max_nodes = max_nodes;
    s->max_devs = max_devs;
    s->max_pods = max_pods;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) s->sm_count = prop.multiProcessorCount;
    s->rank_words = ((max_nodes + 31u) / 32u + gs::kTileWords - 1) / gs::kTileWords * gs::kTileWords;

#define CU_NEW(call)                         \
    do {                                     \
        cudaError_t e_ = (call);             \
        if (e_ != cudaSuccess) {             \
            int rc_ = cuda_fail(e_, #call);  \
            gs_snapshot_destroy(s);          \
            return rc_;                      \
        }                                    \
    } while (0)
    CU_NEW(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    CU_NEW(cudaEventCreate(&s->ev0));
    CU_NEW(cudaEventCreate(&s->ev1));
    CU_NEW(cudaMalloc(&s->d_used, (size_t)max_nodes * max_devs * sizeof(u64)));
    CU_NEW(cudaMalloc(&s->d_cap, (size_t)max_nodes * sizeof(u64)));
    CU_NEW(cudaMalloc(&s->d_nodeinfo, (size_t)max_nodes * sizeof(uint32_t)));
    CU_NEW(cudaMalloc(&s->d_pods, (size_t)max_pods * sizeof(gs_pod_record)));
    CU_NEW(cudaMalloc(&s->d_last_ev, (size_t)max_pods * sizeof(int32_t)));
    CU_NEW(cudaMalloc(&s->d_sorted_keys, 2 * (size_t)GS_MAX_BATCH * sizeof(u64)));
    CU_NEW(cudaMalloc(&s->d_perm, 2 * (size_t)GS_MAX_BATCH * sizeof(uint16_t)));
    CU_NEW(cudaMalloc(&s->d_rank_planes, 2 * (size_t)s->rank_words * gs::kRankPlanes * sizeof(uint32_t)));
    CU_NEW(cudaMalloc(&s->d_node_key, 2 * (size_t)s->rank_words * 32 * sizeof(u64)));
    CU_NEW(cudaMalloc(&s->d_node_key32, 2 * (size_t)s->rank_words * 32 * sizeof(uint32_t)));
    CU_NEW(cudaMalloc(&s->d_row_bucket, 2 * (size_t)GS_MAX_BATCH * sizeof(uint16_t)));
    CU_NEW(cudaMalloc(&s->d_pod_hist, 2 * 512 * sizeof(uint32_t)));
    CU_NEW(cudaMemsetAsync(s->d_pod_hist, 0, 2 * 512 * sizeof(uint32_t), s->stream));
    CU_NEW(cudaMalloc(&s->d_epochs, 6 * sizeof(uint32_t)));
    CU_NEW(cudaMemsetAsync(s->d_epochs, 0, 6 * sizeof(uint32_t), s->stream));
    CU_NEW(cudaMalloc(&s->d_sched, 4 * sizeof(uint32_t)));
    CU_NEW(cudaMemsetAsync(s->d_sched, 0, 4 * sizeof(uint32_t), s->stream));
    CU_NEW(cudaMemsetAsync(s->d_used, 0, (size_t)max_nodes * max_devs * sizeof(u64), s->stream));
    CU_NEW(cudaMemsetAsync(s->d_cap, 0, (size_t)max_nodes * sizeof(u64), s->stream));
    CU_NEW(cudaMemsetAsync(s->d_pods, 0, (size_t)max_pods * sizeof(gs_pod_record), s->stream));
    CU_NEW(cudaMemsetAsync(s->d_rank_planes, 0, 2 * (size_t)s->rank_words * gs::kRankPlanes * sizeof(uint32_t), s->stream));
    CU_NEW(gs::launch_fill_i32(s->d_last_ev, -1, max_pods, s->stream));
    CU_NEW(gs::launch_fill_i32(reinterpret_cast<int32_t*>(s->d_nodeinfo), (int32_t)gs::kUnsetNodeinfo, max_nodes, s->stream));
    CU_NEW(cudaStreamSynchronize(s->stream));
#undef CU_NEW
    s->launches += 2;
    *out = s;
    return GS_OK;
}

int gs_snapshot_destroy(gs_snapshot* s) {
    if (!s) return GS_OK;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    cudaFree(s->d_used);
    cudaFree(s->d_cap);
    cudaFree(s->d_nodeinfo);
    cudaFree(s->d_pods);
    cudaFree(s->d_last_ev);
    cudaFree(s->d_sorted_keys);
    cudaFree(s->d_perm);
    cudaFree(s->d_rank_planes);
    cudaFree(s->d_node_key);
    cudaFree(s->d_node_key32);
    cudaFree(s->d_row_bucket);
    cudaFree(s->d_pod_hist);
    cudaFree(s->d_sched);
    cudaFree(s->d_epochs);
    cudaFree(s->d_stage);
    cudaFree(s->d_bitmap);
    if (s->h_stage) cudaFreeHost(s->h_stage);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->stream) cudaStreamDestroy(s->stream);
    cudaGetLastError();
    delete s;
    return GS_OK;
}

int gs_node_upsert(gs_snapshot* s, const gs_node* nodes, uint32_t n) {
    if (!s || (!nodes && n)) return fail(GS_ERR_INVALID, "null argument");
    Guard g(s);
    int rc = use_device(s);
    if (rc) return rc;
    if (n == 0) return GS_OK;
    uint32_t lo = UINT32_MAX, hi = 0;
    bool contiguous = true;
    for (uint32_t i = 0; i < n; ++i) {
        if (nodes[i].node_idx != nodes[0].node_idx + i) contiguous = false;
        if (nodes[i].node_idx >= s->max_nodes)
            return fail(GS_ERR_CAPACITY, "node_idx %u >= max_nodes %u", nodes[i].node_idx, s->max_nodes);
        if (nodes[i].dev_count > s->max_devs)
            return fail(GS_ERR_CAPACITY, "node %u has %u devices, snapshot stride is %u", nodes[i].node_idx,
                        nodes[i].dev_count, s->max_devs);
        if (nodes[i].node_class > GS_NODE_NOT_SHARE) return fail(GS_ERR_INVALID, "bad node_class");
        lo = std::min(lo, nodes[i].node_idx);
        hi = std::max(hi, nodes[i].node_idx);
        if (nodes[i].node_class == GS_NODE_OK && nodes[i].dev_count > 0 && nodes[i].dev_capacity < (1ull << 40))
            s->max_capacity = std::max<unsigned long long>(s->max_capacity, nodes[i].dev_capacity);
    }
    // Read-modify-write of the covered index range through pinned staging: two contiguous copies
    // each way whatever the order of `nodes` (bulk upserts are contiguous; single ones are tiny).
    const uint32_t span = hi - lo + 1;
    const size_t cap_bytes = (size_t)span * sizeof(u64), info_bytes = (size_t)span * sizeof(uint32_t);
    rc = ensure_host_stage(s, cap_bytes + info_bytes);
    if (rc) return rc;
    CU(cudaStreamSynchronize(s->stream));
    u64* h_cap = reinterpret_cast<u64*>(s->h_stage);
    uint32_t* h_info = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(s->h_stage) + cap_bytes);
    if (!contiguous) {
        CU(cudaMemcpyAsync(h_cap, s->d_cap + lo, cap_bytes, cudaMemcpyDeviceToHost, s->stream));
        CU(cudaMemcpyAsync(h_info, s->d_nodeinfo + lo, info_bytes, cudaMemcpyDeviceToHost, s->stream));
        CU(cudaStreamSynchronize(s->stream));
    }
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t k = nodes[i].node_idx - lo;
        h_cap[k] = nodes[i].dev_capacity;
        h_info[k] = gs::pack_nodeinfo(nodes[i].dev_count, nodes[i].node_class);
    }
    CU(cudaMemcpyAsync(s->d_cap + lo, h_cap, cap_bytes, cudaMemcpyHostToDevice, s->stream));
    CU(cudaMemcpyAsync(s->d_nodeinfo + lo, h_info, info_bytes, cudaMemcpyHostToDevice, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    s->n_nodes = std::max(s->n_nodes, hi + 1);
    return GS_OK;
}

int gs_node_count(gs_snapshot* s, uint32_t* out_nodes) {
    if (!s || !out_nodes) return fail(GS_ERR_INVALID, "null argument");
    Guard g(s);
    *out_nodes = s->n_nodes;
    return GS_OK;
}

int gs_pods_load(gs_snapshot* s, const gs_pod_record* recs, uint32_t first_slot, uint32_t n) {
    if (!s || (!recs && n)) return fail(GS_ERR_INVALID, "null argument");
    Guard g(s);
    int rc = use_device(s);
    if (rc) return rc;
    if ((uint64_t)first_slot + n > s->max_pods) return fail(GS_ERR_CAPACITY, "slots beyond max_pods %u", s->max_pods);
    if (n == 0) return GS_OK;
    rc = h2d(s, s->d_pods + first_slot, recs, (size_t)n * sizeof(gs_pod_record));
    if (rc) return rc;
    rc = timed(s, GS_K_REBUILD, s->stream, 1, [&] {
        return gs::launch_validate_records(s->d_pods, first_slot, n, s->d_nodeinfo, s->n_nodes, s->stream);
    });
    if (rc) return rc;
    CU(cudaStreamSynchronize(s->stream));
    s->pods_hw = std::max(s->pods_hw, first_slot + n);
    return GS_OK;
}

int gs_rebuild(gs_snapshot* s) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    Guard g(s);
    int rc = use_device(s);
    if (rc) return rc;
    rc = timed(s, GS_K_REBUILD, s->stream, 1, [&]() -> cudaError_t {
        cudaError_t e = cudaMemsetAsync(s->d_used, 0, (size_t)s->max_nodes * s->max_devs * sizeof(u64), s->stream);
        if (e != cudaSuccess) return e;
        return gs::launch_rebuild(s->d_pods, s->pods_hw, s->max_devs, s->d_used, s->sm_count, s->stream);
    });
    if (rc) return rc;
    CU(cudaStreamSynchronize(s->stream));
    return GS_OK;
}

int gs_pods_apply(gs_snapshot* s, const gs_pod_event* events, uint32_t n) {
    if (!s || (!events && n)) return fail(GS_ERR_INVALID, "null argument");
    Guard g(s);
    int rc = use_device(s);
    if (rc) return rc;
    if (n == 0) return GS_OK;
    if (n > (1u << 30)) return fail(GS_ERR_INVALID, "too many events in one batch");
    uint32_t hw = s->pods_hw;
    for (uint32_t i = 0; i < n; ++i) {
        if (events[i].slot >= s->max_pods) return fail(GS_ERR_CAPACITY, "event %u: slot %u >= max_pods %u", i, events[i].slot, s->max_pods);
        if (events[i].op > GS_EV_REMOVE) return fail(GS_ERR_INVALID, "event %u: bad op", i);
        hw = std::max(hw, events[i].slot + 1);
    }
    rc = ensure_dev_stage(s, (size_t)n * sizeof(gs_pod_event));
    if (rc) return rc;
    rc = h2d(s, s->d_stage, events, (size_t)n * sizeof(gs_pod_event));
    if (rc) return rc;
    rc = timed(s, GS_K_APPLY, s->stream, 2, [&] {
        return gs::launch_apply_events(reinterpret_cast<const gs_pod_event*>(s->d_stage), n, s->d_pods, s->max_pods,
                                       s->d_last_ev, s->d_nodeinfo, s->n_nodes, s->max_devs, s->d_used, s->stream);
    });
    if (rc) return rc;
    CU(cudaStreamSynchronize(s->stream));
    s->pods_hw = hw;
    return GS_OK;
}

int gs_filter_batch_device(gs_snapshot* s, const uint64_t* d_req, uint32_t n_pods, uint32_t* d_out_bitmap,
                           uint32_t row_stride_words, void* cuda_stream) {
    if (!s || ((!d_req || !d_out_bitmap) && n_pods)) return fail(GS_ERR_INVALID, "null argument");
    Guard g(s, true);
    int rc = use_device(s);
    if (rc) return rc;
    cudaStream_t st = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : s->stream;
    return filter_device_locked(s, reinterpret_cast<const u64*>(d_req), n_pods, d_out_bitmap, row_stride_words, st);
}

int gs_filter_stages_device(gs_snapshot* s, uint32_t stage_mask, const uint64_t* d_req, uint32_t n_pods,
                            uint32_t* d_out_bitmap, uint32_t row_stride_words, void* cuda_stream) {
    if (!s || !d_req || !d_out_bitmap) return fail(GS_ERR_INVALID, "null argument");
    if (n_pods > GS_MAX_BATCH) return fail(GS_ERR_INVALID, "n_pods > GS_MAX_BATCH");
    Guard g(s, true);
    int rc = use_device(s);
    if (rc) return rc;
    if (row_stride_words < bitmap_words(s)) return fail(GS_ERR_INVALID, "row_stride_words too small");
    cudaStream_t st = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : s->stream;
    return filter_chunk(s, reinterpret_cast<const u64*>(d_req), n_pods, d_out_bitmap, row_stride_words, st, stage_mask);
}

int gs_filter_batch(gs_snapshot* s, const uint64_t* req, uint32_t n_pods, uint32_t* out_bitmap,
                    uint32_t row_stride_words) {
    if (!s || ((!req || !out_bitmap) && n_pods)) return fail(GS_ERR_INVALID, "null argument");
    Guard g(s, true);
    int rc = use_device(s);
    if (rc) return rc;
    if (n_pods == 0) return GS_OK;
    if (row_stride_words < bitmap_words(s))
        return fail(GS_ERR_INVALID, "row_stride_words %u < %u", row_stride_words, bitmap_words(s));
    const size_t req_bytes = (size_t)n_pods * sizeof(u64);
    const size_t out_bytes = (size_t)n_pods * row_stride_words * sizeof(uint32_t);
    rc = ensure_dev_stage(s, req_bytes);
    if (rc) return rc;
    rc = ensure_bitmap(s, out_bytes);
    if (rc) return rc;
    rc = h2d(s, s->d_stage, req, req_bytes);
    if (rc) return rc;
    // words beyond the bitmap (caller's padding) are defined: zero
    if (row_stride_words > ((bitmap_words(s) + 3u) & ~3u)) CU(cudaMemsetAsync(s->d_bitmap, 0, out_bytes, s->stream));
    rc = filter_device_locked(s, reinterpret_cast<const u64*>(s->d_stage), n_pods,
                              reinterpret_cast<uint32_t*>(s->d_bitmap), row_stride_words, s->stream);
    if (rc) return rc;
    return d2h(s, out_bitmap, s->d_bitmap, out_bytes);
}

int gs_node_classes(gs_snapshot* s, uint8_t* out_class, uint32_t n) {
    if (!s || (!out_class && n)) return fail(GS_ERR_INVALID, "null argument");
    Guard g(s);
    int rc = use_device(s);
    if (rc) return rc;
    if (n > s->max_nodes) return fail(GS_ERR_CAPACITY, "n > max_nodes");
    if (n == 0) return GS_OK;
    std::vector<uint32_t> info(n);
    rc = d2h(s, info.data(), s->d_nodeinfo, (size_t)n * sizeof(uint32_t));
    if (rc) return rc;
    for (uint32_t i = 0; i < n; ++i) out_class[i] = (uint8_t)gs::nodeinfo_class(info[i]);
    return GS_OK;
}

int gs_bind_batch(gs_snapshot* s, const gs_bind_req* binds, uint32_t n, int32_t* out_dev) {
    if (!s || ((!binds || !out_dev) && n)) return fail(GS_ERR_INVALID, "null argument");
    Guard g(s);
    int rc = use_device(s);
    if (rc) return rc;
    uint32_t hw = s->pods_hw;
    for (uint32_t i = 0; i < n; ++i) {
        if (binds[i].slot >= s->max_pods) return fail(GS_ERR_CAPACITY, "bind %u: slot %u >= max_pods %u", i, binds[i].slot, s->max_pods);
        hw = std::max(hw, binds[i].slot + 1);
    }
    for (uint32_t b0 = 0; b0 < n; b0 += GS_MAX_BATCH) {
        const uint32_t m = std::min<uint32_t>(GS_MAX_BATCH, n - b0);
        const size_t in_bytes = (size_t)m * sizeof(gs_bind_req);
        const size_t out_off = (in_bytes + 255) & ~(size_t)255;
        rc = ensure_dev_stage(s, out_off + (size_t)m * sizeof(int32_t));
        if (rc) return rc;
        gs_bind_req* d_binds = reinterpret_cast<gs_bind_req*>(s->d_stage);
        int32_t* d_out = reinterpret_cast<int32_t*>(reinterpret_cast<char*>(s->d_stage) + out_off);
        rc = h2d(s, d_binds, binds + b0, in_bytes);
        if (rc) return rc;
        rc = timed(s, GS_K_BIND_SORT, s->stream, 1, [&] {
            return gs::launch_sort_binds(s->device, d_binds, m, s->d_sorted_keys, s->d_perm, s->stream);
        });
        if (rc) return rc;
        rc = timed(s, GS_K_BIND_SCAN, s->stream, 1, [&] {
            return gs::launch_bind_scan(d_binds, m, s->d_sorted_keys, s->d_perm, s->d_pods, s->max_pods, s->d_nodeinfo,
                                        s->n_nodes, s->d_cap, s->max_devs, s->d_used, d_out, s->stream);
        });
        if (rc) return rc;
        rc = d2h(s, out_dev + b0, d_out, (size_t)m * sizeof(int32_t));
        if (rc) return rc;
    }
    s->pods_hw = hw;
    return GS_OK;
}

int gs_bind_rollback(gs_snapshot* s, const uint32_t* slots, uint32_t n) {
    if (!s || (!slots && n)) return fail(GS_ERR_INVALID, "null argument");
    std::vector<gs_pod_event> ev(n);
    for (uint32_t i = 0; i < n; ++i) {
        memset(&ev[i], 0, sizeof(gs_pod_event));
        ev[i].slot = slots[i];
        ev[i].op = GS_EV_REMOVE;
    }
    return gs_pods_apply(s, ev.data(), n);
}

int gs_inspect(gs_snapshot* s, uint32_t node_idx, uint64_t* out_capacity, uint64_t* out_used, uint32_t cap_devs,
               uint32_t* out_dev_count, uint8_t* out_class) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    Guard g(s);
    int rc = use_device(s);
    if (rc) return rc;
    if (node_idx >= s->max_nodes) return fail(GS_ERR_CAPACITY, "node_idx %u >= max_nodes %u", node_idx, s->max_nodes);
    uint32_t info = 0;
    u64 cap = 0;
    std::vector<u64> row(s->max_devs);
    rc = d2h(s, &info, s->d_nodeinfo + node_idx, sizeof info);
    if (rc) return rc;
    rc = d2h(s, &cap, s->d_cap + node_idx, sizeof cap);
    if (rc) return rc;
    rc = d2h(s, row.data(), s->d_used + (size_t)node_idx * s->max_devs, (size_t)s->max_devs * sizeof(u64));
    if (rc) return rc;
    const uint32_t ndev = gs::nodeinfo_ndev(info);
    if (out_dev_count) *out_dev_count = ndev;
    if (out_class) *out_class = (uint8_t)gs::nodeinfo_class(info);
    for (uint32_t d = 0; d < ndev && d < cap_devs; ++d) {
        if (out_capacity) out_capacity[d] = cap;
        if (out_used) out_used[d] = row[d];
    }
    return GS_OK;
}

int gs_read_used(gs_snapshot* s, uint32_t first_node, uint32_t n, uint64_t* out_used) {
    if (!s || (!out_used && n)) return fail(GS_ERR_INVALID, "null argument");
    Guard g(s);
    int rc = use_device(s);
    if (rc) return rc;
    if ((uint64_t)first_node + n > s->max_nodes) return fail(GS_ERR_CAPACITY, "nodes beyond max_nodes");
    return d2h(s, out_used, s->d_used + (size_t)first_node * s->max_devs, (size_t)n * s->max_devs * sizeof(u64));
}

int gs_read_pods(gs_snapshot* s, uint32_t first_slot, uint32_t n, gs_pod_record* out) {
    if (!s || (!out && n)) return fail(GS_ERR_INVALID, "null argument");
    Guard g(s);
    int rc = use_device(s);
    if (rc) return rc;
    if ((uint64_t)first_slot + n > s->max_pods) return fail(GS_ERR_CAPACITY, "slots beyond max_pods");
    return d2h(s, out, s->d_pods + first_slot, (size_t)n * sizeof(gs_pod_record));
}

int gs_host_alloc(size_t bytes, void** out) {
    if (!out) return fail(GS_ERR_INVALID, "null out");
    *out = nullptr;
    CU(cudaMallocHost(out, bytes ? bytes : 1));
    return GS_OK;
}
int gs_host_free(void* p) {
    if (!p) return GS_OK;
    CU(cudaFreeHost(p));
    return GS_OK;
}

int gs_set_option(gs_snapshot* s, int option, int value) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    Guard g(s);
    if (option == GS_OPT_PIPELINE_FILTERS) {
        s->pipeline = value != 0;
#ifdef GS_EXPAND_TRACE
        s->pipeline_dev = value == 2;
#endif
        s->prev_filter = false;
        return GS_OK;
    }
    return fail(GS_ERR_INVALID, "unknown option %d", option);
}

int gs_profile_enable(gs_snapshot* s, int on) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    Guard g(s);
    s->profiling = on != 0;
    return GS_OK;
}
int gs_profile_read(gs_snapshot* s, int kernel_id, uint64_t* out_launches, double* out_total_ms) {
    if (!s || kernel_id < 0 || kernel_id >= GS_K_COUNT_) return fail(GS_ERR_INVALID, "bad kernel id");
    Guard g(s);
    if (out_launches) *out_launches = s->prof[kernel_id].launches;
    if (out_total_ms) *out_total_ms = s->prof[kernel_id].total_ms;
    return GS_OK;
}
int gs_profile_reset(gs_snapshot* s) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    Guard g(s);
    for (auto& p : s->prof) p = gs::ProfSlot();
    return GS_OK;
}
int gs_launch_count(gs_snapshot* s, uint64_t* out_launches) {
    if (!s || !out_launches) return fail(GS_ERR_INVALID, "null argument");
    Guard g(s);
    *out_launches = s->launches;
    return GS_OK;
}
int gs_device_ptrs(gs_snapshot* s, void** d_used, void** d_capacity, void** d_nodeinfo, void** d_pods) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    Guard g(s);
    if (d_used) *d_used = s->d_used;
    if (d_capacity) *d_capacity = s->d_cap;
    if (d_nodeinfo) *d_nodeinfo = s->d_nodeinfo;
    if (d_pods) *d_pods = s->d_pods;
    return GS_OK;
}

}  // extern "C"
