// gs_pods_apply_static: the pod-event kernel (K2) for callers that must not allocate.
//
// Who cannot allocate: a scheduler plug-in that has handed all GPU memory management to one owner (the
// host process pins its arena at start-up and forbids later cudaMalloc so that a bind or filter can
// never fail or stall on the allocator), and code running on a latency-critical path where the
// implicit device synchronisation of cudaMalloc / cudaFree is not acceptable.  Such a caller brings
// every buffer itself -- events, pod table, node table and used[] all live in device memory it already
// owns -- and the one piece of scratch the kernel needs (the "last event per slot" table) lives in a
// fixed array inside this library instead of in memory obtained at run time.
#include <stdarg.h>
#include <stdio.h>

#include <mutex>

#include "common.cuh"
#include "state_device.cuh"

namespace {

// The scratch: one int32 per pod slot, -1 = "no event of this batch names the slot".  It is part of the
// library image: every GPU that loads the library carries these 4 MiB until the process exits.
static __device__ __align__(256) int32_t g_last_event[GS_STATIC_MAX_SLOTS];

// Host code may not pass `g_last_event` to a kernel: in host code that name is only a handle for the
// runtime.  The real device address is looked up once per device with cudaGetSymbolAddress.
std::mutex g_setup_mu;
int32_t* g_scratch_addr[64] = {};

int static_fail(int code, const char* fmt, ...) {
    char msg[320];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(msg, sizeof msg, fmt, ap);
    va_end(ap);
    gs::set_last_error(msg);  // read back with gs_last_error()
    return code;
}

__global__ void k_static_clear(int32_t* last_event, uint32_t n_slots) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_slots) last_event[i] = -1;
}

// The two passes are the ones gs_pods_apply runs (state_device.cuh): what an event means is defined once.
// pass 1: the last effective event of every slot
__global__ void k_static_mark(const gs_pod_event* __restrict__ ev, uint32_t n, uint32_t n_slots, int32_t* last_event,
                              const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes, uint32_t dev_stride) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) gs::event_mark(i, ev, n_slots, last_event, nodeinfo, n_nodes, dev_stride);
}

// pass 2: apply exactly those (the scratch is not re-armed here: every call clears what it is about to use)
__global__ void k_static_apply(const gs_pod_event* __restrict__ ev, uint32_t n, gs_pod_record* pods, uint32_t n_slots,
                               int32_t* last_event, const uint32_t* __restrict__ nodeinfo, uint32_t n_nodes,
                               uint32_t dev_stride, unsigned long long* used) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) gs::event_apply<false>(i, ev, pods, n_slots, last_event, nodeinfo, n_nodes, dev_stride, used);
}

}  // namespace

extern "C" {

int gs_pods_apply_static(int device, const gs_pod_event* d_events, uint32_t n_events, gs_pod_record* d_pods,
                         uint32_t n_slots, const uint32_t* d_nodeinfo, uint32_t n_nodes, uint32_t dev_stride,
                         uint64_t* d_used, void* cuda_stream) {
    if (device < 0 || device >= 64) return static_fail(GS_ERR_INVALID, "device %d out of range", device);
    if (n_slots > GS_STATIC_MAX_SLOTS)
        return static_fail(GS_ERR_CAPACITY,
                           "gs_pods_apply_static handles at most %u pod slots (its scratch is a fixed array); got %u -- "
                           "use gs_pods_apply, which allocates its scratch, for larger tables",
                           (unsigned)GS_STATIC_MAX_SLOTS, n_slots);
    if (n_events && (!d_events || !d_pods || !d_nodeinfo || !d_used)) return static_fail(GS_ERR_INVALID, "null device pointer");
    if (dev_stride == 0 || dev_stride > GS_MAX_DEVS) return static_fail(GS_ERR_INVALID, "dev_stride must be 1..%d", GS_MAX_DEVS);
    if (n_events > (1u << 30)) return static_fail(GS_ERR_INVALID, "too many events in one batch");
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return static_fail(GS_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
    int32_t* scratch = nullptr;
    {
        // set-up, once per device, with that device current
        std::lock_guard<std::mutex> lk(g_setup_mu);
        if (!g_scratch_addr[device]) {
            int32_t* p = nullptr;
            e = cudaGetSymbolAddress((void**)&p, g_last_event);
            if (e != cudaSuccess || !p)
                return static_fail(GS_ERR_CUDA, "cudaGetSymbolAddress(g_last_event): %s", cudaGetErrorString(e));
            g_scratch_addr[device] = p;
        }
        scratch = g_scratch_addr[device];
    }
    if (n_events == 0) return GS_OK;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(cuda_stream);
    // the array still holds whatever the previous call (or an interrupted one) left: start from "no event"
    if (n_slots) k_static_clear<<<(n_slots + 255) / 256, 256, 0, st>>>(scratch, n_slots);
    const uint32_t blocks = (n_events + 255) / 256;
    k_static_mark<<<blocks, 256, 0, st>>>(d_events, n_events, n_slots, scratch, d_nodeinfo, n_nodes, dev_stride);
    k_static_apply<<<blocks, 256, 0, st>>>(d_events, n_events, d_pods, n_slots, scratch, d_nodeinfo, n_nodes, dev_stride,
                                           reinterpret_cast<unsigned long long*>(d_used));
    e = cudaGetLastError();
    if (e != cudaSuccess) return static_fail(GS_ERR_CUDA, "kernel launch: %s", cudaGetErrorString(e));
    return GS_OK;
}

}  // extern "C"
