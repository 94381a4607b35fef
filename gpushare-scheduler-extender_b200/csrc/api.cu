// C-ABI layer (include/gpushare_b200.h) over the snapshot object and the kernel launchers.
//
// Locking follows the reference's cache (pkg/cache/nodeinfo.go:116,141): Filter calls share the snapshot
// (each runs in its own filter context: stream, staging and K3 work buffers), everything that changes the
// cluster state or reads it back is exclusive.  Work that an entry point leaves running on the GPU is
// ordered against later calls on other streams with events, recorded only when a later call needs them.
#include <cuda.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <mutex>
#include <new>
#include <vector>

#include "common.cuh"
#include "kernels.h"

using u64 = unsigned long long;

static thread_local char g_err[640] = "";

static int fail(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
static int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}
static const char* fault_name(unsigned int w) {
    switch (w) {
        case 1: return "k_prep waited for the expand kernel that last read its work buffers";
        case 2: return "k_prep pod blocks waited for each other's bucket counts";
        case 3: return "k_expand waited for its batch's k_prep";
        case 4: return "k_expand waited for a bulk copy of tile inputs";
        case 5: return "k_expand waited for a bulk copy of node keys";
        default: return "unknown wait";
    }
}
// `s` may be null (calls that have no snapshot)
static int cuda_fail(const gs_snapshot* s, cudaError_t e, const char* what) {
    const unsigned int w = (s && s->h_fault) ? *reinterpret_cast<volatile unsigned int*>(s->h_fault) : 0u;
    if (w)
        return fail(GS_ERR_CUDA, "%s: %s (%s); a kernel gave up a bounded wait: %s", what, cudaGetErrorName(e),
                    cudaGetErrorString(e), fault_name(w));
    return fail(GS_ERR_CUDA, "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
}
#define CU(call)                                               \
    do {                                                       \
        cudaError_t e_ = (call);                               \
        if (e_ != cudaSuccess) return cuda_fail(s, e_, #call); \
    } while (0)

namespace gs {
void set_last_error(const char* msg) { snprintf(g_err, sizeof g_err, "%s", msg); }
}  // namespace gs

namespace {

using gs::FilterCtx;

int use_device(gs_snapshot* s) {
    CU(cudaSetDevice(s->device));
    return GS_OK;
}

// SMs the calling context can use: the device's, or fewer inside a green context / an MPS partition.
// (Driver entry points are looked up at run time so the library does not link against libcuda.)
int context_sm_count(int device) {
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) sms = 148;
    typedef CUresult (*get_current_t)(CUcontext*);
    typedef CUresult (*get_res_t)(CUcontext, CUdevResource*, CUdevResourceType);
    void *f_cur = nullptr, *f_res = nullptr;
    cudaDriverEntryPointQueryResult q1, q2;
    if (cudaGetDriverEntryPoint("cuCtxGetCurrent", &f_cur, cudaEnableDefault, &q1) == cudaSuccess &&
        cudaGetDriverEntryPoint("cuCtxGetDevResource", &f_res, cudaEnableDefault, &q2) == cudaSuccess &&
        q1 == cudaDriverEntryPointSuccess && q2 == cudaDriverEntryPointSuccess && f_cur && f_res) {
        CUcontext cur = nullptr;
        if (reinterpret_cast<get_current_t>(f_cur)(&cur) == CUDA_SUCCESS && cur) {
            CUdevResource res;
            memset(&res, 0, sizeof res);
            if (reinterpret_cast<get_res_t>(f_res)(cur, &res, CU_DEV_RESOURCE_TYPE_SM) == CUDA_SUCCESS &&
                res.sm.smCount > 0 && (int)res.sm.smCount < sms)
                sms = (int)res.sm.smCount;
        }
    }
    cudaGetLastError();
    return sms > 0 ? sms : 148;
}

int grow_device(gs_snapshot* s, void** p, size_t* have, size_t bytes, bool doubling) {
    if (bytes <= *have) return GS_OK;
    const size_t want = doubling ? std::max(bytes, *have * 2) : bytes;
    if (*p) CU(cudaFree(*p));
    *p = nullptr;
    *have = 0;
    CU(cudaMalloc(p, want));
    *have = want;
    return GS_OK;
}
int grow_pinned(gs_snapshot* s, void** p, size_t* have, size_t bytes) {
    if (bytes <= *have) return GS_OK;
    const size_t want = std::max(bytes, *have * 2);
    if (*p) CU(cudaFreeHost(*p));
    *p = nullptr;
    *have = 0;
    CU(cudaMallocHost(p, want));
    *have = want;
    return GS_OK;
}

// Blocks handed out by gs_host_alloc: known to be page-locked and (unified addressing) addressable by kernels under the
// same pointer, without asking the driver on every call.
std::mutex g_host_mu;
std::map<uintptr_t, size_t> g_host_blocks;  // start -> bytes
bool in_host_block(const void* p) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(p);
    std::lock_guard<std::mutex> lk(g_host_mu);
    auto it = g_host_blocks.upper_bound(a);
    if (it == g_host_blocks.begin()) return false;
    --it;
    return a < it->first + it->second;
}

// Page-locked host memory that kernels of this device can address (cudaHostAlloc / cudaHostRegister under unified
// addressing): the pointer kernels should use, or nullptr.
void* device_view_of_host(const void* p) {
    if (in_host_block(p)) return const_cast<void*>(p);
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
}

bool is_pinned(const void* p) {
    if (in_host_block(p)) return true;
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

// Host -> device copy of a caller buffer that may be pageable: pageable data is first copied into pinned
// staging so the DMA runs at full rate and asynchronously.  `h_stage` is the staging of whoever owns `st`.
int h2d(gs_snapshot* s, cudaStream_t st, void** h_stage, size_t* h_bytes, void* dst, const void* src, size_t bytes) {
    if (bytes == 0) return GS_OK;
    if (is_pinned(src)) {
        CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st));
        return GS_OK;
    }
    CU(cudaStreamSynchronize(st));  // staging may still be in flight from an earlier copy
    int rc = grow_pinned(s, h_stage, h_bytes, bytes);
    if (rc) return rc;
    memcpy(*h_stage, src, bytes);
    CU(cudaMemcpyAsync(dst, *h_stage, bytes, cudaMemcpyHostToDevice, st));
    return GS_OK;
}
// Device -> host into a caller buffer; ends synchronised.
int d2h(gs_snapshot* s, cudaStream_t st, void** h_stage, size_t* h_bytes, void* dst, const void* src, size_t bytes) {
    if (bytes == 0) return GS_OK;
    if (is_pinned(dst)) {
        CU(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        return GS_OK;
    }
    CU(cudaStreamSynchronize(st));
    int rc = grow_pinned(s, h_stage, h_bytes, bytes);
    if (rc) return rc;
    CU(cudaMemcpyAsync(*h_stage, src, bytes, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    memcpy(dst, *h_stage, bytes);
    return GS_OK;
}
int w_h2d(gs_snapshot* s, void* dst, const void* src, size_t bytes) {
    return h2d(s, s->stream, &s->h_stage, &s->h_stage_bytes, dst, src, bytes);
}
int w_d2h(gs_snapshot* s, void* dst, const void* src, size_t bytes) {
    return d2h(s, s->stream, &s->h_stage, &s->h_stage_bytes, dst, src, bytes);
}

// Runs one launcher, optionally bracketed by events for the per-kernel counters.
template <class F>
int timed(gs_snapshot* s, int kid, cudaStream_t st, uint64_t n_launches, F&& f) {
    if (!s->profiling) {
        cudaError_t e = f();
        if (e != cudaSuccess) return cuda_fail(s, e, "kernel launch");
        s->launches += n_launches;
        return GS_OK;
    }
    std::lock_guard<std::mutex> lk(s->prof_mu);
    CU(cudaEventRecord(s->ev0, st));
    cudaError_t e = f();
    if (e != cudaSuccess) return cuda_fail(s, e, "kernel launch");
    s->launches += n_launches;
    CU(cudaEventRecord(s->ev1, st));
    CU(cudaEventSynchronize(s->ev1));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
    s->prof[kid].launches += n_launches;
    s->prof[kid].total_ms += ms;
    return GS_OK;
}

uint32_t bitmap_words(const gs_snapshot* s) { return (s->n_nodes + 31u) / 32u; }

// ---- locks -------------------------------------------------------------------------------------
using ReadLock = std::shared_lock<gs::RwLock>;
using WriteLock = std::unique_lock<gs::RwLock>;

// A writer goes behind whatever filters left running on other streams, and ends their pipelined chains.
// (Called with the exclusive lock held: no filter call is inside the library.)
int begin_write(gs_snapshot* s) {
    int rc = use_device(s);
    if (rc) return rc;
    for (FilterCtx& c : s->ctx) {
        c.prev_filter = false;
        if (c.outstanding) {
            if (c.last_stream != s->stream) {
                CU(cudaEventRecord(c.done, c.last_stream));
                CU(cudaStreamWaitEvent(s->stream, c.done, 0));
            }
            c.outstanding = false;
        }
    }
    return GS_OK;
}
// An asynchronous write was enqueued on the writers' stream: later filters on other streams go behind it.
int end_async_write(gs_snapshot* s) {
    CU(cudaEventRecord(s->ev_write, s->stream));
    ++s->write_seq;
    return GS_OK;
}

struct CtxLease {
    FilterCtx* c = nullptr;
    std::unique_lock<std::mutex> lk;
};
// Context 0 serves the device-buffer calls (its pipelining state follows one stream of calls); host-buffer
// calls take whichever of the others is free, and queue behind one of them when all are busy.
CtxLease lease_ctx(gs_snapshot* s, bool device_call) {
    CtxLease l;
    if (device_call) {
        l.c = &s->ctx[0];
        l.lk = std::unique_lock<std::mutex>(l.c->mu);
        return l;
    }
    // the lowest free one first: a single caller keeps using the context whose staging buffers it has already grown
    for (int k = 1; k < gs::kFilterCtxs; ++k) {
        FilterCtx* c = &s->ctx[k];
        std::unique_lock<std::mutex> lk(c->mu, std::try_to_lock);
        if (lk.owns_lock()) {
            l.c = c;
            l.lk = std::move(lk);
            return l;
        }
    }
    const uint32_t start = s->ctx_rr.fetch_add(1, std::memory_order_relaxed);
    l.c = &s->ctx[1 + start % (gs::kFilterCtxs - 1)];
    l.lk = std::unique_lock<std::mutex>(l.c->mu);
    return l;
}

// Before a context enqueues on `st`: behind the last asynchronous write, and behind the context's own earlier
// work if that went to another stream (the work buffers are shared between the calls of a context).
int order_filter(gs_snapshot* s, FilterCtx& c, cudaStream_t st) {
    if (c.outstanding && c.last_stream != st) {
        CU(cudaEventRecord(c.done, c.last_stream));
        CU(cudaStreamWaitEvent(st, c.done, 0));
        c.prev_filter = false;
        c.seen_write_seq = ~0ull;  // this stream has not been ordered behind anything yet
    }
    if (c.seen_write_seq != s->write_seq || c.last_stream != st) {
        if (s->write_seq != 0 && st != s->stream) CU(cudaStreamWaitEvent(st, s->ev_write, 0));
        c.seen_write_seq = s->write_seq;
        c.prev_filter = false;
    }
    c.last_stream = st;
    return GS_OK;
}

int alloc_ctx(gs_snapshot* s, FilterCtx& c) {
    const size_t B = GS_MAX_BATCH;
    CU(cudaStreamCreateWithFlags(&c.stream, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&c.done, cudaEventDisableTiming));
    CU(cudaMalloc(&c.d_grouped_req, 2 * B * sizeof(u64)));
    CU(cudaMalloc(&c.d_perm, 2 * B * sizeof(uint16_t)));
    CU(cudaMalloc(&c.d_row_bucket, 2 * B * sizeof(uint16_t)));
    CU(cudaMalloc(&c.d_pod_hist, 2 * 512 * sizeof(uint32_t)));
    CU(cudaMalloc(&c.d_rank_planes, 2 * (size_t)s->rank_words * gs::kRankPlanes * sizeof(uint32_t)));
    CU(cudaMalloc(&c.d_node_key, 2 * (size_t)s->rank_words * 32 * sizeof(u64)));
    CU(cudaMalloc(&c.d_node_key32, 2 * (size_t)s->rank_words * 32 * sizeof(uint32_t)));
    CU(cudaMalloc(&c.d_sched, 4 * sizeof(uint32_t)));
    CU(cudaMalloc(&c.d_epochs, 8 * sizeof(uint32_t)));
    CU(cudaMalloc(&c.d_cslices, (B + 1024) * sizeof(u64)));
    CU(cudaMalloc(&c.d_csorted, B * sizeof(u64)));
    CU(cudaMalloc(&c.d_cdup, B * sizeof(uint32_t)));
    CU(cudaMalloc(&c.d_creq, B * sizeof(u64)));
    CU(cudaMalloc(&c.d_ccut, B * sizeof(uint16_t)));
    CU(cudaMemsetAsync(c.d_pod_hist, 0, 2 * 512 * sizeof(uint32_t), s->stream));
    CU(cudaMemsetAsync(c.d_sched, 0, 4 * sizeof(uint32_t), s->stream));
    CU(cudaMemsetAsync(c.d_epochs, 0, 8 * sizeof(uint32_t), s->stream));
    CU(cudaMemsetAsync(c.d_cdup, 0, B * sizeof(uint32_t), s->stream));
    CU(cudaMemsetAsync(c.d_rank_planes, 0, 2 * (size_t)s->rank_words * gs::kRankPlanes * sizeof(uint32_t), s->stream));
    return GS_OK;
}
void free_ctx(FilterCtx& c) {
    if (c.stream) cudaStreamSynchronize(c.stream);
    cudaFree(c.d_grouped_req);
    cudaFree(c.d_perm);
    cudaFree(c.d_row_bucket);
    cudaFree(c.d_pod_hist);
    cudaFree(c.d_rank_planes);
    cudaFree(c.d_node_key);
    cudaFree(c.d_node_key32);
    cudaFree(c.d_sched);
    cudaFree(c.d_epochs);
    cudaFree(c.d_cslices);
    cudaFree(c.d_csorted);
    cudaFree(c.d_cdup);
    cudaFree(c.d_creq);
    cudaFree(c.d_ccut);
    cudaFree(c.d_stage);
    cudaFree(c.d_bitmap);
    if (c.h_stage) cudaFreeHost(c.h_stage);
    if (c.done) cudaEventDestroy(c.done);
    if (c.stream) cudaStreamDestroy(c.stream);
}

// The two filter kernels for one chunk of <= GS_MAX_BATCH pods; everything on `st`.
int filter_chunk(gs_snapshot* s, FilterCtx& c, const u64* d_req, uint32_t n_pods, uint32_t* d_out, uint32_t row_stride,
                 cudaStream_t st, uint32_t stages = 7u) {
    const uint32_t words = bitmap_words(s);
    if (n_pods == 0 || words == 0) return GS_OK;
    const uint32_t words_pad = (words + gs::kTileWords - 1) / gs::kTileWords * gs::kTileWords;
    if (words_pad > s->rank_words) return fail(GS_ERR_CAPACITY, "rank planes too small");
    const uint32_t set = c.fb_set;
    // full launches alternate; single-stage profiling launches stay on one set (development builds can let
    // expand-only launches alternate and overlap too: GS_EXPAND_TRACE + option value 2)
    const bool dev_overlap_expand = s->pipeline_dev && stages == GS_STAGE_EXPAND;
    if (stages == 7u || dev_overlap_expand) c.fb_set ^= 1u;
    gs::FilterBuffers fb;
    fb.grouped_req = c.d_grouped_req + (size_t)set * GS_MAX_BATCH;
    fb.perm = c.d_perm + (size_t)set * GS_MAX_BATCH;
    fb.row_bucket = c.d_row_bucket + (size_t)set * GS_MAX_BATCH;
    fb.pod_hist = c.d_pod_hist + set * 512;
    fb.planes = c.d_rank_planes + (size_t)set * s->rank_words * gs::kRankPlanes;
    fb.node_key = c.d_node_key + (size_t)set * s->rank_words * 32;
    fb.node_key32 = c.d_node_key32 + (size_t)set * s->rank_words * 32;
    fb.sched = c.d_sched + set * 2;
    fb.set_index = set;
    fb.prep_done = c.d_epochs + 3 * set;
    fb.prep_epoch = c.d_epochs + 3 * set + 1;
    fb.expand_epoch = c.d_epochs + 3 * set + 2;
    fb.fault = s->d_fault;
    fb.early_starts = c.d_epochs + 6;
    // With GS_OPT_PIPELINE_FILTERS, a filter that directly follows another filter of this context on the same
    // stream, writing somewhere else, does not wait for that filter's expand kernel to drain before its own prep
    // kernel starts (the two touch different work-buffer sets).  Anything else keeps plain stream order.
    const char* out_lo = reinterpret_cast<const char*>(d_out);
    const char* out_hi = out_lo + (size_t)n_pods * row_stride * sizeof(uint32_t);
    const bool overlap_ok = dev_overlap_expand ||
                            (s->pipeline && stages == 7u && c.prev_filter && c.prev_stream == st &&
                             (out_hi <= c.prev_out_lo || out_lo >= c.prev_out_hi));
    // bucket scale: requests up to the largest device capacity spread over all 512 buckets
    uint32_t bits = 0;
    while (bits < 64 && (s->max_capacity >> bits) != 0) ++bits;
    const uint32_t shift = bits > 9 ? bits - 9 : 0;
    int rc = GS_OK;
    const bool do_pods = (stages & GS_STAGE_SORT) != 0, do_nodes = (stages & GS_STAGE_NODE_RANK) != 0;
    if (do_pods || do_nodes)
        rc = timed(s, do_pods ? GS_K_SORT : GS_K_NODE_RANK, st, 1, [&] {
            cudaError_t e = gs::launch_prep(d_req, n_pods, do_pods, do_nodes, shift, fb, s->d_used, s->d_cap, s->d_nodeinfo,
                                            s->n_nodes, s->max_devs, words_pad, overlap_ok, c.n_expand[set], &c.n_prep_blocks[set], st);
            if (e == cudaSuccess) ++c.n_prep[set];
            return e;
        });
    if (rc) return rc;
    if (stages & GS_STAGE_EXPAND)
        rc = timed(s, GS_K_EXPAND, st, 1, [&] {
            cudaError_t e = gs::launch_expand(s->sm_count, fb, n_pods, words, row_stride, d_out, overlap_ok, c.n_prep[set], c.n_prep_blocks[set], st);
            if (e == cudaSuccess) ++c.n_expand[set];
            return e;
        });
    c.prev_filter = rc == GS_OK && stages == 7u && !s->profiling;
    c.prev_stream = st;
    c.prev_out_lo = out_lo;
    c.prev_out_hi = out_hi;
    return rc;
}

int filter_device_locked(gs_snapshot* s, FilterCtx& c, const u64* d_req, uint32_t n_pods, uint32_t* d_out,
                         uint32_t row_stride, cudaStream_t st) {
    if (row_stride < bitmap_words(s)) return fail(GS_ERR_INVALID, "row_stride_words %u < %u", row_stride, bitmap_words(s));
    for (uint32_t p0 = 0; p0 < n_pods; p0 += GS_MAX_BATCH) {
        const uint32_t n = std::min<uint32_t>(GS_MAX_BATCH, n_pods - p0);
        int rc = filter_chunk(s, c, d_req + p0, n, d_out + (size_t)p0 * row_stride, row_stride, st);
        if (rc) return rc;
    }
    return GS_OK;
}

// words per row k_expand really writes for this stride and alignment (see launch_expand)
uint32_t expand_words_out(uint32_t words, uint32_t row_stride, const void* d_out) {
    const bool vec = (row_stride % 4 == 0) && ((reinterpret_cast<uintptr_t>(d_out) & 15u) == 0);
    if (!vec) return words;
    uint32_t w = (words + 7u) & ~7u;
    if (w > row_stride) w = (words + 3u) & ~3u;
    return w;
}

// host_req / host_cut: d_req / d_cutoff are the device's view of page-locked host memory
int compact_device_locked(gs_snapshot* s, FilterCtx& c, const u64* d_req, uint32_t n_pods, uint16_t* d_rank,
                          uint32_t rank_stride, uint16_t* d_cutoff, cudaStream_t st, bool host_req = false, bool host_cut = false) {
    if (rank_stride < s->n_nodes) return fail(GS_ERR_INVALID, "rank_stride %u < node count %u", rank_stride, s->n_nodes);
    c.prev_filter = false;  // (a snapshot without nodes still answers: the cutoffs depend on the requests only)
    for (uint32_t p0 = 0, k = 0; p0 < n_pods; p0 += GS_MAX_BATCH, ++k) {
        const uint32_t n = std::min<uint32_t>(GS_MAX_BATCH, n_pods - p0);
        uint32_t launched = 0;
        cudaError_t le = cudaSuccess;
        int rc = timed(s, GS_K_COMPACT, st, 0, [&] {
            gs::RankDsts dsts = {};
            dsts.p[0] = d_rank + (size_t)k * rank_stride;
            dsts.n = 1;
            dsts.offset = 0;
            dsts.count = s->n_nodes;
            le = gs::launch_compact(d_req + p0, n, c.d_cslices, c.d_csorted, c.d_cdup, s->d_used, s->d_cap, s->d_nodeinfo,
                                    s->n_nodes, s->max_devs, dsts, d_cutoff + p0, s->sm_count, st, &launched,
                                    host_req ? c.d_creq : nullptr, host_cut ? c.d_ccut : nullptr);
            return le;
        });
        s->launches += launched;
        if (rc) return rc;
    }
    return GS_OK;
}

}  // namespace

extern "C" {

const char* gs_last_error(void) { return g_err; }
int gs_abi_version(void) { return GS_ABI_VERSION; }

int gs_device_count(int* out_count) {
    const gs_snapshot* s = nullptr;
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
    {
        const gs_snapshot* s = nullptr;
        CU(cudaSetDevice(device));
        CU(cudaFree(0));  // the context exists from here on
    }
    gs_snapshot* s = new (std::nothrow) gs_snapshot();
    if (!s) return fail(GS_ERR_NOMEM, "out of host memory");
    s->device = device;
    s->max_nodes = max_nodes;
    s->max_devs = max_devs;
    s->max_pods = max_pods;
    s->sm_count = context_sm_count(device);
    s->rank_words = ((max_nodes + 31u) / 32u + gs::kTileWords - 1) / gs::kTileWords * gs::kTileWords;

#define CU_NEW(call)                           \
    do {                                       \
        cudaError_t e_ = (call);               \
        if (e_ != cudaSuccess) {               \
            int rc_ = cuda_fail(s, e_, #call); \
            gs_snapshot_destroy(s);            \
            return rc_;                        \
        }                                      \
    } while (0)
    CU_NEW(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    CU_NEW(cudaEventCreate(&s->ev0));
    CU_NEW(cudaEventCreate(&s->ev1));
    CU_NEW(cudaEventCreateWithFlags(&s->ev_write, cudaEventDisableTiming));
    for (int k = 0; k < gs::kApplyRing; ++k) CU_NEW(cudaEventCreateWithFlags(&s->ring_free[k], cudaEventDisableTiming));
    CU_NEW(cudaHostAlloc(reinterpret_cast<void**>(&s->h_fault), sizeof(unsigned int), cudaHostAllocMapped));
    *s->h_fault = 0;
    CU_NEW(cudaHostGetDevicePointer(reinterpret_cast<void**>(&s->d_fault), s->h_fault, 0));
    CU_NEW(cudaMalloc(&s->d_used, (size_t)max_nodes * max_devs * sizeof(u64)));
    CU_NEW(cudaMalloc(&s->d_cap, (size_t)max_nodes * sizeof(u64)));
    CU_NEW(cudaMalloc(&s->d_nodeinfo, (size_t)max_nodes * sizeof(uint32_t)));
    CU_NEW(cudaMalloc(&s->d_pods, (size_t)max_pods * sizeof(gs_pod_record)));
    CU_NEW(cudaMalloc(&s->d_last_ev, (size_t)max_pods * sizeof(int32_t)));
    CU_NEW(cudaMalloc(&s->d_bind_keys, (size_t)GS_MAX_BATCH * sizeof(u64)));
    CU_NEW(cudaMalloc(&s->d_bind_perm, (size_t)GS_MAX_BATCH * sizeof(uint16_t)));
    CU_NEW(cudaMemsetAsync(s->d_used, 0, (size_t)max_nodes * max_devs * sizeof(u64), s->stream));
    CU_NEW(cudaMemsetAsync(s->d_cap, 0, (size_t)max_nodes * sizeof(u64), s->stream));
    CU_NEW(cudaMemsetAsync(s->d_pods, 0, (size_t)max_pods * sizeof(gs_pod_record), s->stream));
    CU_NEW(gs::launch_fill_i32(s->d_last_ev, -1, max_pods, s->stream));
    CU_NEW(gs::launch_fill_i32(reinterpret_cast<int32_t*>(s->d_nodeinfo), (int32_t)gs::kUnsetNodeinfo, max_nodes, s->stream));
    for (FilterCtx& c : s->ctx) {
        int rc = alloc_ctx(s, c);
        if (rc) {
            gs_snapshot_destroy(s);
            return rc;
        }
    }
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
    for (FilterCtx& c : s->ctx) free_ctx(c);
    cudaFree(s->d_used);
    cudaFree(s->d_cap);
    cudaFree(s->d_nodeinfo);
    cudaFree(s->d_pods);
    cudaFree(s->d_last_ev);
    cudaFree(s->d_bind_keys);
    cudaFree(s->d_bind_perm);
    cudaFree(s->d_stage);
    if (s->h_stage) cudaFreeHost(s->h_stage);
    for (int k = 0; k < gs::kApplyRing; ++k) {
        cudaFree(s->d_ring[k]);
        if (s->h_ring[k]) cudaFreeHost(s->h_ring[k]);
        if (s->ring_free[k]) cudaEventDestroy(s->ring_free[k]);
    }
    if (s->h_fault) cudaFreeHost(s->h_fault);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->ev_write) cudaEventDestroy(s->ev_write);
    if (s->stream) cudaStreamDestroy(s->stream);
    cudaGetLastError();
    delete s;
    return GS_OK;
}

int gs_node_upsert(gs_snapshot* s, const gs_node* nodes, uint32_t n) {
    if (!s || (!nodes && n)) return fail(GS_ERR_INVALID, "null argument");
    WriteLock lk(s->state_mu);
    int rc = begin_write(s);
    if (rc) return rc;
    if (n == 0) return GS_OK;
    uint32_t lo = UINT32_MAX, hi = 0;
    bool contiguous = true;
    unsigned long long max_cap = s->max_capacity;
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
            max_cap = std::max<unsigned long long>(max_cap, nodes[i].dev_capacity);
    }
    // Read-modify-write of the covered index range through pinned staging: two contiguous copies
    // each way whatever the order of `nodes` (bulk upserts are contiguous; single ones are tiny).
    const uint32_t span = hi - lo + 1;
    const size_t cap_bytes = (size_t)span * sizeof(u64), info_bytes = (size_t)span * sizeof(uint32_t);
    CU(cudaStreamSynchronize(s->stream));
    rc = grow_pinned(s, &s->h_stage, &s->h_stage_bytes, cap_bytes + info_bytes);
    if (rc) return rc;
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
    s->max_capacity = max_cap;
    return GS_OK;
}

int gs_node_count(gs_snapshot* s, uint32_t* out_nodes) {
    if (!s || !out_nodes) return fail(GS_ERR_INVALID, "null argument");
    ReadLock lk(s->state_mu);
    *out_nodes = s->n_nodes;
    return GS_OK;
}

int gs_pods_load(gs_snapshot* s, const gs_pod_record* recs, uint32_t first_slot, uint32_t n) {
    if (!s || (!recs && n)) return fail(GS_ERR_INVALID, "null argument");
    WriteLock lk(s->state_mu);
    int rc = begin_write(s);
    if (rc) return rc;
    if ((uint64_t)first_slot + n > s->max_pods) return fail(GS_ERR_CAPACITY, "slots beyond max_pods %u", s->max_pods);
    if (n == 0) return GS_OK;
    rc = w_h2d(s, s->d_pods + first_slot, recs, (size_t)n * sizeof(gs_pod_record));
    if (rc) return rc;
    rc = timed(s, GS_K_REBUILD, s->stream, 1, [&] {
        return gs::launch_validate_records(s->d_pods, first_slot, n, s->d_nodeinfo, s->n_nodes, s->max_devs, s->stream);
    });
    if (rc) return rc;
    CU(cudaStreamSynchronize(s->stream));
    return GS_OK;
}

int gs_rebuild(gs_snapshot* s) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    WriteLock lk(s->state_mu);
    int rc = begin_write(s);
    if (rc) return rc;
    // every slot of the table is read (slots never written are zero = invalid), so records that reached the
    // table by any road -- gs_pods_load, events, binds, gs_pods_apply_static on the snapshot's arrays -- count
    rc = timed(s, GS_K_REBUILD, s->stream, 2, [&]() -> cudaError_t {
        return gs::launch_rebuild(s->d_pods, s->max_pods, s->max_nodes, s->max_devs, s->d_used, s->sm_count, s->stream);
    });
    if (rc) return rc;
    CU(cudaStreamSynchronize(s->stream));
    return GS_OK;
}

static int check_events(gs_snapshot* s, const gs_pod_event* events, uint32_t n) {
    if (n > (1u << 30)) return fail(GS_ERR_INVALID, "too many events in one batch");
    for (uint32_t i = 0; i < n; ++i) {
        if (events[i].slot >= s->max_pods)
            return fail(GS_ERR_CAPACITY, "event %u: slot %u >= max_pods %u", i, events[i].slot, s->max_pods);
        if (events[i].op > GS_EV_REMOVE) return fail(GS_ERR_INVALID, "event %u: bad op", i);
    }
    return GS_OK;
}

int gs_pods_apply(gs_snapshot* s, const gs_pod_event* events, uint32_t n) {
    if (!s || (!events && n)) return fail(GS_ERR_INVALID, "null argument");
    WriteLock lk(s->state_mu);
    int rc = begin_write(s);
    if (rc) return rc;
    if (n == 0) return GS_OK;
    rc = check_events(s, events, n);
    if (rc) return rc;
    const size_t ev_bytes = (size_t)n * sizeof(gs_pod_event);
    rc = grow_device(s, &s->d_stage, &s->d_stage_bytes, ev_bytes, true);
    if (rc) return rc;
    rc = grow_pinned(s, &s->h_stage, &s->h_stage_bytes, ev_bytes);
    if (rc) return rc;
    memcpy(s->h_stage, events, ev_bytes);  // (writers end synchronised: the staging is free when a call starts)
    CU(cudaMemcpyAsync(s->d_stage, s->h_stage, ev_bytes, cudaMemcpyHostToDevice, s->stream));
    rc = timed(s, GS_K_APPLY, s->stream, 2, [&] {
        return gs::launch_apply_events(reinterpret_cast<const gs_pod_event*>(s->d_stage), n, s->d_pods, s->max_pods,
                                       s->d_last_ev, s->d_nodeinfo, s->n_nodes, s->max_devs, s->d_used, s->stream);
    });
    if (rc) return rc;
    CU(cudaStreamSynchronize(s->stream));
    return GS_OK;
}

int gs_pods_apply_async(gs_snapshot* s, const gs_pod_event* events, uint32_t n) {
    if (!s || (!events && n)) return fail(GS_ERR_INVALID, "null argument");
    WriteLock lk(s->state_mu);
    int rc = begin_write(s);
    if (rc) return rc;
    if (n == 0) return GS_OK;
    rc = check_events(s, events, n);
    if (rc) return rc;
    // the events are copied into one of a few pinned slots, so the caller's array is free when the call
    // returns; a slot is reused once the copy that read it has run
    const uint32_t k = s->ring_next++ % gs::kApplyRing;
    if (s->ring_used[k]) CU(cudaEventSynchronize(s->ring_free[k]));
    const size_t bytes = (size_t)n * sizeof(gs_pod_event);
    if (bytes > s->ring_bytes[k]) {
        const size_t want = std::max(bytes, s->ring_bytes[k] * 2);
        if (s->ring_used[k]) CU(cudaStreamSynchronize(s->stream));  // the kernels that read the old device slot
        if (s->h_ring[k]) CU(cudaFreeHost(s->h_ring[k]));
        if (s->d_ring[k]) CU(cudaFree(s->d_ring[k]));
        s->h_ring[k] = s->d_ring[k] = nullptr;
        s->ring_bytes[k] = 0;
        CU(cudaMallocHost(&s->h_ring[k], want));
        CU(cudaMalloc(&s->d_ring[k], want));
        s->ring_bytes[k] = want;
    }
    memcpy(s->h_ring[k], events, bytes);
    CU(cudaMemcpyAsync(s->d_ring[k], s->h_ring[k], bytes, cudaMemcpyHostToDevice, s->stream));
    rc = timed(s, GS_K_APPLY, s->stream, 2, [&] {
        return gs::launch_apply_events(reinterpret_cast<const gs_pod_event*>(s->d_ring[k]), n, s->d_pods, s->max_pods,
                                       s->d_last_ev, s->d_nodeinfo, s->n_nodes, s->max_devs, s->d_used, s->stream);
    });
    if (rc) return rc;
    CU(cudaEventRecord(s->ring_free[k], s->stream));
    s->ring_used[k] = true;
    return end_async_write(s);
}

int gs_snapshot_sync(gs_snapshot* s) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    WriteLock lk(s->state_mu);
    int rc = begin_write(s);
    if (rc) return rc;
    CU(cudaStreamSynchronize(s->stream));
    return GS_OK;
}

int gs_snapshot_external_write(gs_snapshot* s, void* cuda_stream) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    WriteLock lk(s->state_mu);
    int rc = begin_write(s);
    if (rc) return rc;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    if (st != s->stream) {
        // the writers' stream goes behind the caller's work; everything else is ordered behind the writers' stream
        cudaEvent_t ev = nullptr;
        CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        cudaError_t e = cudaEventRecord(ev, st);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(s->stream, ev, 0);
        cudaEventDestroy(ev);
        if (e != cudaSuccess) return cuda_fail(s, e, "ordering behind the caller's stream");
    }
    return end_async_write(s);
}

int gs_filter_batch_device(gs_snapshot* s, const uint64_t* d_req, uint32_t n_pods, uint32_t* d_out_bitmap,
                           uint32_t row_stride_words, void* cuda_stream) {
    if (!s || ((!d_req || !d_out_bitmap) && n_pods)) return fail(GS_ERR_INVALID, "null argument");
    ReadLock lk(s->state_mu);
    CtxLease l = lease_ctx(s, true);
    int rc = use_device(s);
    if (rc) return rc;
    if (n_pods == 0) return GS_OK;
    cudaStream_t st = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : s->stream;
    rc = order_filter(s, *l.c, st);
    if (rc) return rc;
    l.c->outstanding = true;
    return filter_device_locked(s, *l.c, reinterpret_cast<const u64*>(d_req), n_pods, d_out_bitmap, row_stride_words, st);
}

int gs_filter_stages_device(gs_snapshot* s, uint32_t stage_mask, const uint64_t* d_req, uint32_t n_pods,
                            uint32_t* d_out_bitmap, uint32_t row_stride_words, void* cuda_stream) {
    if (!s || !d_req || !d_out_bitmap) return fail(GS_ERR_INVALID, "null argument");
    if (n_pods > GS_MAX_BATCH) return fail(GS_ERR_INVALID, "n_pods > GS_MAX_BATCH");
    ReadLock lk(s->state_mu);
    CtxLease l = lease_ctx(s, true);
    int rc = use_device(s);
    if (rc) return rc;
    if (row_stride_words < bitmap_words(s)) return fail(GS_ERR_INVALID, "row_stride_words too small");
    cudaStream_t st = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : s->stream;
    rc = order_filter(s, *l.c, st);
    if (rc) return rc;
    l.c->outstanding = true;
    return filter_chunk(s, *l.c, reinterpret_cast<const u64*>(d_req), n_pods, d_out_bitmap, row_stride_words, st, stage_mask);
}

int gs_filter_batch(gs_snapshot* s, const uint64_t* req, uint32_t n_pods, uint32_t* out_bitmap,
                    uint32_t row_stride_words) {
    if (!s || ((!req || !out_bitmap) && n_pods)) return fail(GS_ERR_INVALID, "null argument");
    ReadLock lk(s->state_mu);
    CtxLease l = lease_ctx(s, false);
    FilterCtx& c = *l.c;
    int rc = use_device(s);
    if (rc) return rc;
    if (n_pods == 0) return GS_OK;
    const uint32_t words = bitmap_words(s);
    if (row_stride_words < words) return fail(GS_ERR_INVALID, "row_stride_words %u < %u", row_stride_words, words);
    const size_t req_bytes = (size_t)n_pods * sizeof(u64);
    const size_t out_bytes = (size_t)n_pods * row_stride_words * sizeof(uint32_t);
    rc = grow_device(s, &c.d_stage, &c.d_stage_bytes, req_bytes, true);
    if (rc) return rc;
    rc = grow_device(s, &c.d_bitmap, &c.d_bitmap_bytes, out_bytes, false);
    if (rc) return rc;
    rc = order_filter(s, c, c.stream);
    if (rc) return rc;
    rc = h2d(s, c.stream, &c.h_stage, &c.h_stage_bytes, c.d_stage, req, req_bytes);
    if (rc) return rc;
    // every word of the caller's rows is defined: what the kernel does not write is zero
    if (row_stride_words > expand_words_out(words, row_stride_words, c.d_bitmap) || words == 0)
        CU(cudaMemsetAsync(c.d_bitmap, 0, out_bytes, c.stream));
    rc = filter_device_locked(s, c, reinterpret_cast<const u64*>(c.d_stage), n_pods, reinterpret_cast<uint32_t*>(c.d_bitmap),
                              row_stride_words, c.stream);
    if (rc) return rc;
    return d2h(s, c.stream, &c.h_stage, &c.h_stage_bytes, out_bitmap, c.d_bitmap, out_bytes);
}

int gs_filter_batch_subset(gs_snapshot* s, const uint64_t* req, uint32_t n_pods, const uint32_t* node_subset,
                           uint32_t n_candidates, uint32_t* out_bitmap, uint32_t row_stride_words) {
    if (!s || ((!req || !out_bitmap) && n_pods) || (!node_subset && n_candidates))
        return fail(GS_ERR_INVALID, "null argument");
    ReadLock lk(s->state_mu);
    CtxLease l = lease_ctx(s, false);
    FilterCtx& c = *l.c;
    int rc = use_device(s);
    if (rc) return rc;
    const uint32_t words = (n_candidates + 31u) / 32u;
    if (row_stride_words < words) return fail(GS_ERR_INVALID, "row_stride_words %u < %u", row_stride_words, words);
    if (n_pods == 0) return GS_OK;
    const size_t out_bytes = (size_t)n_pods * row_stride_words * sizeof(uint32_t);
    if (n_candidates == 0) {
        memset(out_bitmap, 0, out_bytes);
        return GS_OK;
    }
    const size_t req_bytes = ((size_t)n_pods * sizeof(u64) + 255) & ~(size_t)255;
    const size_t sub_bytes = (size_t)n_candidates * sizeof(uint32_t);
    // A small request -- what kube-scheduler sends: one pod, a list of candidate names -- goes through the context's
    // page-locked block alone: inputs copied there by the CPU, the kernel reads them and writes the answer in place over
    // PCIe, one wait, no DMA set up at all (three copies cost more than the kernel: 38 -> ~15 us per call).
    {
        const size_t sub_pad = (sub_bytes + 255) & ~(size_t)255;
        const size_t total = req_bytes + sub_pad + out_bytes;
        if (n_pods <= 64 && total <= (64u << 10) && !s->profiling) {
            rc = grow_pinned(s, &c.h_stage, &c.h_stage_bytes, total);
            if (rc) return rc;
            rc = order_filter(s, c, c.stream);
            if (rc) return rc;
            c.prev_filter = false;
            char* h = static_cast<char*>(c.h_stage);
            memcpy(h, req, (size_t)n_pods * sizeof(u64));
            memcpy(h + req_bytes, node_subset, sub_bytes);
            if (row_stride_words > words) memset(h + req_bytes + sub_pad, 0, out_bytes);
            cudaError_t e = gs::launch_filter_subset(reinterpret_cast<const u64*>(h), n_pods,
                                                     reinterpret_cast<const uint32_t*>(h + req_bytes), n_candidates, s->d_used,
                                                     s->d_cap, s->d_nodeinfo, s->n_nodes, s->max_devs,
                                                     reinterpret_cast<uint32_t*>(h + req_bytes + sub_pad), row_stride_words, c.stream);
            if (e != cudaSuccess) return cuda_fail(s, e, "kernel launch");
            s->launches += 1;
            CU(cudaStreamSynchronize(c.stream));
            memcpy(out_bitmap, h + req_bytes + sub_pad, out_bytes);
            return GS_OK;
        }
    }
    rc = grow_device(s, &c.d_stage, &c.d_stage_bytes, req_bytes + sub_bytes, true);
    if (rc) return rc;
    rc = grow_device(s, &c.d_bitmap, &c.d_bitmap_bytes, out_bytes, false);
    if (rc) return rc;
    rc = order_filter(s, c, c.stream);
    if (rc) return rc;
    c.prev_filter = false;
    char* stage = reinterpret_cast<char*>(c.d_stage);
    // two small copies; when either source is pageable both go through the pinned staging one after the other
    rc = h2d(s, c.stream, &c.h_stage, &c.h_stage_bytes, stage, req, (size_t)n_pods * sizeof(u64));
    if (rc) return rc;
    rc = h2d(s, c.stream, &c.h_stage, &c.h_stage_bytes, stage + req_bytes, node_subset, sub_bytes);
    if (rc) return rc;
    if (row_stride_words > words) CU(cudaMemsetAsync(c.d_bitmap, 0, out_bytes, c.stream));
    rc = timed(s, GS_K_SUBSET, c.stream, 1, [&] {
        return gs::launch_filter_subset(reinterpret_cast<const u64*>(stage), n_pods,
                                        reinterpret_cast<const uint32_t*>(stage + req_bytes), n_candidates, s->d_used, s->d_cap,
                                        s->d_nodeinfo, s->n_nodes, s->max_devs, reinterpret_cast<uint32_t*>(c.d_bitmap),
                                        row_stride_words, c.stream);
    });
    if (rc) return rc;
    return d2h(s, c.stream, &c.h_stage, &c.h_stage_bytes, out_bitmap, c.d_bitmap, out_bytes);
}

int gs_filter_batch_compact_device(gs_snapshot* s, const uint64_t* d_req, uint32_t n_pods, uint16_t* d_out_rank,
                                   uint32_t rank_stride, uint16_t* d_out_cutoff, void* cuda_stream) {
    if (!s || ((!d_req || !d_out_rank || !d_out_cutoff) && n_pods)) return fail(GS_ERR_INVALID, "null argument");
    ReadLock lk(s->state_mu);
    CtxLease l = lease_ctx(s, true);
    int rc = use_device(s);
    if (rc) return rc;
    if (n_pods == 0) return GS_OK;
    cudaStream_t st = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : s->stream;
    rc = order_filter(s, *l.c, st);
    if (rc) return rc;
    l.c->outstanding = true;
    // (buffers from gs_host_alloc are "device" buffers too: the kernels then touch host memory as little as they can)
    return compact_device_locked(s, *l.c, reinterpret_cast<const u64*>(d_req), n_pods, d_out_rank, rank_stride, d_out_cutoff, st,
                                 in_host_block(d_req), in_host_block(d_out_cutoff));
}

int gs_filter_batch_compact_device_gather(gs_snapshot* s, const uint64_t* d_req, uint32_t n_pods, uint16_t* const* d_rank_dsts,
                                          uint32_t n_dsts, uint32_t dst_offset, uint32_t slice_count, uint16_t* d_out_cutoff,
                                          void* cuda_stream) {
    if (!s || !d_req || !d_rank_dsts || !d_out_cutoff) return fail(GS_ERR_INVALID, "null argument");
    if (n_dsts == 0 || n_dsts > (uint32_t)gs::kMaxRankDsts) return fail(GS_ERR_INVALID, "n_dsts must be 1..%d", gs::kMaxRankDsts);
    if (n_pods == 0 || n_pods > GS_MAX_BATCH) return fail(GS_ERR_INVALID, "n_pods must be 1..GS_MAX_BATCH (ranks are relative to one chunk)");
    ReadLock lk(s->state_mu);
    CtxLease l = lease_ctx(s, true);
    FilterCtx& c = *l.c;
    int rc = use_device(s);
    if (rc) return rc;
    if (slice_count < s->n_nodes) return fail(GS_ERR_INVALID, "slice_count %u < node count %u", slice_count, s->n_nodes);
    cudaStream_t st = cuda_stream ? reinterpret_cast<cudaStream_t>(cuda_stream) : s->stream;
    rc = order_filter(s, c, st);
    if (rc) return rc;
    c.outstanding = true;
    c.prev_filter = false;
    gs::RankDsts dsts = {};
    for (uint32_t k = 0; k < n_dsts; ++k) {
        if (!d_rank_dsts[k]) return fail(GS_ERR_INVALID, "null destination %u", k);
        dsts.p[k] = d_rank_dsts[k];
    }
    dsts.n = n_dsts;
    dsts.offset = dst_offset;
    dsts.count = slice_count;
    uint32_t launched = 0;
    rc = timed(s, GS_K_COMPACT, st, 0, [&] {
        return gs::launch_compact(reinterpret_cast<const u64*>(d_req), n_pods, c.d_cslices, c.d_csorted, c.d_cdup, s->d_used, s->d_cap,
                                  s->d_nodeinfo, s->n_nodes, s->max_devs, dsts, d_out_cutoff, s->sm_count, st, &launched);
    });
    s->launches += launched;
    return rc;
}

int gs_filter_batch_compact(gs_snapshot* s, const uint64_t* req, uint32_t n_pods, uint16_t* out_rank, uint32_t rank_stride,
                            uint16_t* out_cutoff) {
    if (!s || ((!req || !out_rank || !out_cutoff) && n_pods)) return fail(GS_ERR_INVALID, "null argument");
    ReadLock lk(s->state_mu);
    CtxLease l = lease_ctx(s, false);
    FilterCtx& c = *l.c;
    int rc = use_device(s);
    if (rc) return rc;
    if (n_pods == 0) return GS_OK;
    if (rank_stride < s->n_nodes) return fail(GS_ERR_INVALID, "rank_stride %u < node count %u", rank_stride, s->n_nodes);
    const uint32_t n_chunks = (n_pods + GS_MAX_BATCH - 1) / GS_MAX_BATCH;
    // Page-locked caller buffers: the kernels read the requests and write rank[] / cutoff[] in place, over PCIe -- 80 KB
    // in and 120 KB out at config C4's size, less than what three copies cost to set up.  The call still ends
    // synchronised: the results are in the caller's memory when it returns.
    if (rank_stride == s->n_nodes && s->n_nodes != 0) {
        const u64* v_req = static_cast<const u64*>(device_view_of_host(req));
        uint16_t* v_rank = static_cast<uint16_t*>(device_view_of_host(out_rank));
        uint16_t* v_cut = static_cast<uint16_t*>(device_view_of_host(out_cutoff));
        if (v_req && v_rank && v_cut) {
            rc = order_filter(s, c, c.stream);
            if (rc) return rc;
            c.outstanding = true;
            rc = compact_device_locked(s, c, v_req, n_pods, v_rank, rank_stride, v_cut, c.stream, true, true);
            if (rc) return rc;
            CU(cudaStreamSynchronize(c.stream));
            return GS_OK;
        }
    }
    const size_t req_bytes = ((size_t)n_pods * sizeof(u64) + 255) & ~(size_t)255;
    const size_t rank_bytes = (size_t)n_chunks * rank_stride * sizeof(uint16_t);
    const size_t rank_pad = (rank_bytes + 255) & ~(size_t)255;
    const size_t cut_bytes = (size_t)n_pods * sizeof(uint16_t);
    rc = grow_device(s, &c.d_stage, &c.d_stage_bytes, req_bytes + rank_pad + cut_bytes, true);
    if (rc) return rc;
    rc = order_filter(s, c, c.stream);
    if (rc) return rc;
    char* stage = reinterpret_cast<char*>(c.d_stage);
    rc = h2d(s, c.stream, &c.h_stage, &c.h_stage_bytes, stage, req, (size_t)n_pods * sizeof(u64));
    if (rc) return rc;
    if (rank_stride > s->n_nodes || s->n_nodes == 0) CU(cudaMemsetAsync(stage + req_bytes, 0, rank_bytes, c.stream));
    rc = compact_device_locked(s, c, reinterpret_cast<const u64*>(stage), n_pods, reinterpret_cast<uint16_t*>(stage + req_bytes),
                               rank_stride, reinterpret_cast<uint16_t*>(stage + req_bytes + rank_pad), c.stream);
    if (rc) return rc;
    if (is_pinned(out_rank) && is_pinned(out_cutoff)) {  // both copies in flight together, one wait
        CU(cudaMemcpyAsync(out_rank, stage + req_bytes, rank_bytes, cudaMemcpyDeviceToHost, c.stream));
        CU(cudaMemcpyAsync(out_cutoff, stage + req_bytes + rank_pad, cut_bytes, cudaMemcpyDeviceToHost, c.stream));
        CU(cudaStreamSynchronize(c.stream));
        return GS_OK;
    }
    rc = d2h(s, c.stream, &c.h_stage, &c.h_stage_bytes, out_rank, stage + req_bytes, rank_bytes);
    if (rc) return rc;
    return d2h(s, c.stream, &c.h_stage, &c.h_stage_bytes, out_cutoff, stage + req_bytes + rank_pad, cut_bytes);
}

int gs_node_classes(gs_snapshot* s, uint8_t* out_class, uint32_t n) {
    if (!s || (!out_class && n)) return fail(GS_ERR_INVALID, "null argument");
    WriteLock lk(s->state_mu);
    int rc = begin_write(s);
    if (rc) return rc;
    if (n > s->max_nodes) return fail(GS_ERR_CAPACITY, "n > max_nodes");
    if (n == 0) return GS_OK;
    std::vector<uint32_t> info(n);
    rc = w_d2h(s, info.data(), s->d_nodeinfo, (size_t)n * sizeof(uint32_t));
    if (rc) return rc;
    for (uint32_t i = 0; i < n; ++i) out_class[i] = (uint8_t)gs::nodeinfo_class(info[i]);
    return GS_OK;
}

#ifndef GS_BIND_IN_PLACE_MAX
#define GS_BIND_IN_PLACE_MAX 1024
#endif
constexpr uint32_t kBindInPlaceMax = GS_BIND_IN_PLACE_MAX;

int gs_bind_batch(gs_snapshot* s, const gs_bind_req* binds, uint32_t n, int32_t* out_dev) {
    if (!s || ((!binds || !out_dev) && n)) return fail(GS_ERR_INVALID, "null argument");
    WriteLock lk(s->state_mu);
    int rc = begin_write(s);
    if (rc) return rc;
    if (n == 0) return GS_OK;
    {
        // A slot may be named once per call: binds of different nodes run in parallel, and two of them filling one
        // slot would race (pkg/cache/nodeinfo.go:197-202 adds one fresh map entry per committed pod).
        std::vector<uint32_t> slots(n);
        for (uint32_t i = 0; i < n; ++i) {
            if (binds[i].slot >= s->max_pods)
                return fail(GS_ERR_CAPACITY, "bind %u: slot %u >= max_pods %u", i, binds[i].slot, s->max_pods);
            slots[i] = binds[i].slot;
        }
        std::sort(slots.begin(), slots.end());
        for (uint32_t i = 1; i < n; ++i)
            if (slots[i] == slots[i - 1]) return fail(GS_ERR_INVALID, "slot %u is named by more than one bind of the batch", slots[i]);
    }
    // One pinned block holds the chunk's binds and its results: copy in, two kernels, copy out, ONE wait.  (Writers are
    // exclusive and end synchronised, so the staging is free when a call starts.)
    for (uint32_t b0 = 0; b0 < n; b0 += GS_MAX_BATCH) {
        const uint32_t m = std::min<uint32_t>(GS_MAX_BATCH, n - b0);
        const size_t in_bytes = (size_t)m * sizeof(gs_bind_req);
        const size_t out_off = (in_bytes + 255) & ~(size_t)255;
        const size_t out_bytes = (size_t)m * sizeof(int32_t);
        rc = grow_device(s, &s->d_stage, &s->d_stage_bytes, out_off + out_bytes, true);
        if (rc) return rc;
        rc = grow_pinned(s, &s->h_stage, &s->h_stage_bytes, out_off + out_bytes);
        if (rc) return rc;
        char* h = reinterpret_cast<char*>(s->h_stage);
        gs_bind_req* d_binds = reinterpret_cast<gs_bind_req*>(s->d_stage);
        int32_t* d_out = reinterpret_cast<int32_t*>(reinterpret_cast<char*>(s->d_stage) + out_off);
        memcpy(h, binds + b0, in_bytes);
        // a handful of binds (the extender binds one pod per request): the kernels read them from the page-locked block
        // and write the chosen devices there, no copies to set up
        const bool in_place = m <= kBindInPlaceMax && !s->profiling;
        if (in_place) {
            d_binds = reinterpret_cast<gs_bind_req*>(h);
            d_out = reinterpret_cast<int32_t*>(h + out_off);
        } else {
            CU(cudaMemcpyAsync(d_binds, h, in_bytes, cudaMemcpyHostToDevice, s->stream));
        }
        if (m <= 32) {  // one launch of one warp orders and walks a short queue
            rc = timed(s, GS_K_BIND_SCAN, s->stream, 1, [&] {
                return gs::launch_bind_small(d_binds, m, s->d_pods, s->max_pods, s->d_nodeinfo, s->n_nodes, s->d_cap, s->max_devs,
                                             s->d_used, d_out, s->stream);
            });
            if (rc) return rc;
        } else {
            rc = timed(s, GS_K_BIND_SORT, s->stream, 1, [&] {
                return gs::launch_sort_binds(s->device, d_binds, m, s->d_bind_keys, s->d_bind_perm, s->stream);
            });
            if (rc) return rc;
            rc = timed(s, GS_K_BIND_SCAN, s->stream, 1, [&] {
                return gs::launch_bind_scan(d_binds, m, s->d_bind_keys, s->d_bind_perm, s->d_pods, s->max_pods, s->d_nodeinfo,
                                            s->n_nodes, s->d_cap, s->max_devs, s->d_used, d_out, s->stream);
            });
            if (rc) return rc;
        }
        if (!in_place) CU(cudaMemcpyAsync(h + out_off, d_out, out_bytes, cudaMemcpyDeviceToHost, s->stream));
        CU(cudaStreamSynchronize(s->stream));
        memcpy(out_dev + b0, h + out_off, out_bytes);
    }
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
    WriteLock lk(s->state_mu);
    int rc = begin_write(s);
    if (rc) return rc;
    if (node_idx >= s->max_nodes) return fail(GS_ERR_CAPACITY, "node_idx %u >= max_nodes %u", node_idx, s->max_nodes);
    // one pinned block, three small copies in flight together, one wait
    const size_t row_bytes = (size_t)s->max_devs * sizeof(u64);
    CU(cudaStreamSynchronize(s->stream));
    rc = grow_pinned(s, &s->h_stage, &s->h_stage_bytes, row_bytes + 16);
    if (rc) return rc;
    char* h = reinterpret_cast<char*>(s->h_stage);
    CU(cudaMemcpyAsync(h, s->d_used + (size_t)node_idx * s->max_devs, row_bytes, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaMemcpyAsync(h + row_bytes, s->d_cap + node_idx, sizeof(u64), cudaMemcpyDeviceToHost, s->stream));
    CU(cudaMemcpyAsync(h + row_bytes + 8, s->d_nodeinfo + node_idx, sizeof(uint32_t), cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    const u64* row = reinterpret_cast<const u64*>(h);
    const u64 cap = *reinterpret_cast<const u64*>(h + row_bytes);
    const uint32_t info = *reinterpret_cast<const uint32_t*>(h + row_bytes + 8);
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
    WriteLock lk(s->state_mu);
    int rc = begin_write(s);
    if (rc) return rc;
    if ((uint64_t)first_node + n > s->max_nodes) return fail(GS_ERR_CAPACITY, "nodes beyond max_nodes");
    return w_d2h(s, out_used, s->d_used + (size_t)first_node * s->max_devs, (size_t)n * s->max_devs * sizeof(u64));
}

int gs_read_pods(gs_snapshot* s, uint32_t first_slot, uint32_t n, gs_pod_record* out) {
    if (!s || (!out && n)) return fail(GS_ERR_INVALID, "null argument");
    WriteLock lk(s->state_mu);
    int rc = begin_write(s);
    if (rc) return rc;
    if ((uint64_t)first_slot + n > s->max_pods) return fail(GS_ERR_CAPACITY, "slots beyond max_pods");
    return w_d2h(s, out, s->d_pods + first_slot, (size_t)n * sizeof(gs_pod_record));
}

int gs_host_alloc(size_t bytes, void** out) {
    const gs_snapshot* s = nullptr;
    if (!out) return fail(GS_ERR_INVALID, "null out");
    *out = nullptr;
    CU(cudaMallocHost(out, bytes ? bytes : 1));
    {
        std::lock_guard<std::mutex> lk(g_host_mu);
        g_host_blocks[reinterpret_cast<uintptr_t>(*out)] = bytes ? bytes : 1;
    }
    return GS_OK;
}
int gs_host_free(void* p) {
    const gs_snapshot* s = nullptr;
    if (!p) return GS_OK;
    {
        std::lock_guard<std::mutex> lk(g_host_mu);
        g_host_blocks.erase(reinterpret_cast<uintptr_t>(p));
    }
    CU(cudaFreeHost(p));
    return GS_OK;
}

int gs_set_option(gs_snapshot* s, int option, int value) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    WriteLock lk(s->state_mu);
    if (option == GS_OPT_PIPELINE_FILTERS) {
        s->pipeline = value != 0;
#ifdef GS_EXPAND_TRACE
        s->pipeline_dev = value == 2;
#endif
        for (FilterCtx& c : s->ctx) c.prev_filter = false;
        return GS_OK;
    }
    return fail(GS_ERR_INVALID, "unknown option %d", option);
}

int gs_profile_enable(gs_snapshot* s, int on) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    WriteLock lk(s->state_mu);
    s->profiling = on != 0;
    for (FilterCtx& c : s->ctx) c.prev_filter = false;
    return GS_OK;
}
int gs_profile_read(gs_snapshot* s, int kernel_id, uint64_t* out_launches, double* out_total_ms) {
    if (!s || kernel_id < 0 || kernel_id >= GS_K_COUNT_) return fail(GS_ERR_INVALID, "bad kernel id");
    std::lock_guard<std::mutex> lk(s->prof_mu);
    if (out_launches) *out_launches = s->prof[kernel_id].launches;
    if (out_total_ms) *out_total_ms = s->prof[kernel_id].total_ms;
    return GS_OK;
}
int gs_profile_reset(gs_snapshot* s) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    std::lock_guard<std::mutex> lk(s->prof_mu);
    for (auto& p : s->prof) p = gs::ProfSlot();
    return GS_OK;
}
int gs_launch_count(gs_snapshot* s, uint64_t* out_launches) {
    if (!s || !out_launches) return fail(GS_ERR_INVALID, "null argument");
    *out_launches = s->launches.load();
    return GS_OK;
}
int gs_sm_count(gs_snapshot* s, int* out_sms) {
    if (!s || !out_sms) return fail(GS_ERR_INVALID, "null argument");
    *out_sms = s->sm_count;
    return GS_OK;
}
int gs_debug_early_starts(gs_snapshot* s, uint64_t* out_count) {
    if (!s || !out_count) return fail(GS_ERR_INVALID, "null argument");
    *out_count = 0;
    WriteLock lk(s->state_mu);  // (nothing else of the snapshot is in a call)
    int rc = use_device(s);
    if (rc) return rc;
    CU(cudaDeviceSynchronize());
    for (FilterCtx& c : s->ctx) {
        uint32_t v = 0;
        if (!c.d_epochs) continue;
        CU(cudaMemcpy(&v, c.d_epochs + 6, sizeof(v), cudaMemcpyDeviceToHost));
        *out_count += v;
    }
    return GS_OK;
}
int gs_device_ptrs(gs_snapshot* s, void** d_used, void** d_capacity, void** d_nodeinfo, void** d_pods) {
    if (!s) return fail(GS_ERR_INVALID, "null snapshot");
    ReadLock lk(s->state_mu);
    if (d_used) *d_used = s->d_used;
    if (d_capacity) *d_capacity = s->d_cap;
    if (d_nodeinfo) *d_nodeinfo = s->d_nodeinfo;
    if (d_pods) *d_pods = s->d_pods;
    return GS_OK;
}

}  // extern "C"
