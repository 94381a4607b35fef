/*
 * gpushare_ext.h -- host-side mirror of the reference's cache / predicate / bind / inspect interface on top
 * of the C ABI of gpushare_b200.h.
 *
 * The reference is Go; no Go toolchain exists where this was built, so the host logic that a cgo shim would
 * carry (name -> node index and (UID, node, device) -> slot tables, annotation parsing, the lazily created
 * and then frozen NodeInfo, reason strings, the JSON envelope of the filter response) is written in C++ here,
 * with the same names, argument meaning and error texts as the reference:
 *
 *   reference                                                          | here
 *   -------------------------------------------------------------------+---------------------------
 *   SchedulerCache.GetNodeInfo        pkg/cache/cache.go:130-162         | gs_ext_get_node_info
 *   SchedulerCache.BuildCache         pkg/cache/cache.go:49-74           | gs_ext_build_cache
 *   SchedulerCache.AddOrUpdatePod     pkg/cache/cache.go:89-113          | gs_ext_add_or_update_pod
 *   SchedulerCache.RemovePod          pkg/cache/cache.go:116-127         | gs_ext_remove_pod
 *   SchedulerCache.KnownPod           pkg/cache/cache.go:81-87           | gs_ext_known_pod
 *   Predicate.Handler + predicate     pkg/scheduler/predicate.go:15-39,  | gs_ext_filter_json,
 *                                     gpushare-predicate.go:16-37        | gs_ext_filter_codes
 *   Bind.Handler + NodeInfo.Allocate  pkg/scheduler/gpushare-bind.go:21-65, pkg/cache/nodeinfo.go:139-206 | gs_ext_bind
 *   Inspect buildNode                 pkg/scheduler/inspect.go:33-71     | gs_ext_inspect_node
 *   informer handlers + syncPod       pkg/gpushare/controller.go:77-100,174-332 | gs_ext_informer_*, gs_ext_ctrl_drain
 *   utils.* parsers                   pkg/utils/pod.go:45-60,94-113,146-155 | gs_ext_pod_* getters
 *
 * All decisions (pass sets, chosen devices, used memory) come from the CUDA kernels; this layer only
 * translates.  Errors: functions return 0 on success, non-zero otherwise, with the reference's message in
 * gs_last_error() (or in the err buffer where one is passed).
 */
#ifndef GPUSHARE_EXT_H
#define GPUSHARE_EXT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* v1.PodPhase */
enum { GS_PHASE_PENDING = 0, GS_PHASE_RUNNING = 1, GS_PHASE_SUCCEEDED = 2, GS_PHASE_FAILED = 3, GS_PHASE_UNKNOWN = 4 };
/* outcome of the predicate for one node, in the order the reference tests them */
enum { GS_FILTER_PASS = 0, GS_FILTER_NOT_FOUND = 1, GS_FILTER_NOT_SHARE = 2, GS_FILTER_INSUFFICIENT = 3 };
/* what the apiserver does during Allocate (the real calls stay in the host language; tests script them) */
enum {
    GS_API_OK = 0,
    GS_API_UPDATE_CONFLICT_THEN_OK = 1,
    GS_API_UPDATE_CONFLICT_GET_FAILS = 2,
    GS_API_UPDATE_CONFLICT_TWICE = 3,
    GS_API_UPDATE_OTHER_ERROR = 4,
    GS_API_BIND_FAILS = 5
};

typedef struct gs_ext gs_ext;
typedef struct gs_ext_pod gs_ext_pod;

/* ---- the slice of v1.Pod the path reads (no GPU needed) ---- */
gs_ext_pod* gs_ext_pod_new(const char* name, const char* ns, const char* uid);
void gs_ext_pod_free(gs_ext_pod* p);
void gs_ext_pod_set_node_name(gs_ext_pod* p, const char* node_name);
void gs_ext_pod_set_phase(gs_ext_pod* p, int phase);
void gs_ext_pod_set_deleting(gs_ext_pod* p, int has_deletion_timestamp);
void gs_ext_pod_set_annotation(gs_ext_pod* p, const char* key, const char* value);
void gs_ext_pod_add_container(gs_ext_pod* p, int has_limit, int64_t limit_value);
/* pkg/utils/pod.go parsers */
int64_t gs_ext_atoi(const char* s, int* err);               /* strconv.Atoi: err 0 nil, 1 syntax, 2 range */
int64_t gs_ext_pod_gpu_id(const gs_ext_pod* p);             /* GetGPUIDFromAnnotation        pod.go:45-60   */
uint64_t gs_ext_pod_annotation_mem(const gs_ext_pod* p);    /* GetGPUMemoryFromPodAnnotation pod.go:94-113  */
int64_t gs_ext_pod_resource_mem(const gs_ext_pod* p);       /* GetGPUMemoryFromPodResource   pod.go:146-155 */

/* ---- the extender state: listers + cache mirror over one gs_snapshot ---- */
int gs_ext_create(int device, uint32_t max_nodes, uint32_t max_devs, uint32_t max_pods, gs_ext** out);
int gs_ext_destroy(gs_ext* x);
/* the gs_snapshot underneath (gpushare_b200.h), e.g. for its launch and profile counters */
struct gs_snapshot* gs_ext_snapshot(gs_ext* x);

int gs_ext_node_lister_set(gs_ext* x, const char* name, int has_mem, int64_t mem, int has_count, int64_t count);
int gs_ext_node_lister_delete(gs_ext* x, const char* name);
int gs_ext_pod_lister_set(gs_ext* x, const gs_ext_pod* p);
int gs_ext_pod_lister_delete(gs_ext* x, const char* ns, const char* name);
/* the stand-in apiserver's pod store, consulted by getPod (gpushare-bind.go:50-62) and the retry of Allocate */
int gs_ext_api_set_pod(gs_ext* x, const gs_ext_pod* p);

int gs_ext_get_node_info(gs_ext* x, const char* name);
int gs_ext_build_cache(gs_ext* x);
int gs_ext_add_or_update_pod(gs_ext* x, const gs_ext_pod* p);
int gs_ext_remove_pod(gs_ext* x, const gs_ext_pod* p);
int gs_ext_known_pod(gs_ext* x, const char* uid);

/* One Filter request -> the HTTP response body the reference would marshal (malloc'd; gs_ext_free). */
char* gs_ext_filter_json(gs_ext* x, const gs_ext_pod* pod, const char* const* node_names, size_t n);
/* Many Filter requests over the same candidate list in ONE device batch; codes is [n_pods][n] GS_FILTER_*. */
int gs_ext_filter_codes(gs_ext* x, const gs_ext_pod* const* pods, size_t n_pods, const char* const* node_names,
                        size_t n, uint8_t* codes);
/* Bind.Handler: returns 0 when ExtenderBindingResult.Error would be "", else 1 with the message in err.
 * *dev_id receives allocateGPUID's choice (-1 = none).  Binds to one node take turns (NodeInfo.rwmu,
 * nodeinfo.go:141), binds to different nodes run side by side, the apiserver step included. */
int gs_ext_bind(gs_ext* x, const char* pod_name, const char* pod_ns, const char* pod_uid, const char* node,
                int api_mode, int64_t* dev_id, char* err, size_t errcap);
/* The apiserver step of Allocate (Pods.Update + Pods.Bind, nodeinfo.go:150-184) is scripted by api_mode; a hook, if
 * set, is called where those round trips happen, with only the node's lock held (tests use it to hold a bind
 * inside the step and show that a bind to another node does not wait). */
typedef void (*gs_ext_api_hook)(void* user, const char* pod_key, const char* node, int api_mode);
int gs_ext_set_api_hook(gs_ext* x, gs_ext_api_hook hook, void* user);
/* Inspect: per-device total/used of a cached node; returns gpu-count, or -1 with err. */
int gs_ext_inspect_node(gs_ext* x, const char* name, uint64_t* total, uint64_t* used, int cap, char* err, size_t errcap);

/* Event path (pkg/gpushare/controller.go:77-100,174-332): informer handlers behind the IsGPUsharingPod filter, the
 * work queue, and syncPod.  Update the pod lister first, as the informer does, then report the event;
 * gs_ext_ctrl_drain runs syncPod for every queued key. */
int gs_ext_informer_add(gs_ext* x, const gs_ext_pod* p);
int gs_ext_informer_update(gs_ext* x, const gs_ext_pod* old_p, const gs_ext_pod* new_p);
int gs_ext_informer_delete(gs_ext* x, const gs_ext_pod* p);
int gs_ext_ctrl_drain(gs_ext* x);
size_t gs_ext_ctrl_queue_len(gs_ext* x);

void gs_ext_free(void* p);

#ifdef __cplusplus
}
#endif
#endif
