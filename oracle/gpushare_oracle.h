/*
 * gpushare_oracle.h -- CPU restatement of the gpushare-scheduler-extender hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it, and
 * there only as the checker or as the timed CPU baseline -- never as the thing shipped.
 *
 * PARITY STATUS: "parity unpinned" by the reference's own tests -- the reference has no tests,
 * fixtures or golden files at all, and neither Go nor the vendored
 * beego/logs package is available, so the reference cannot be compiled or run here.  The oracle is
 * pinned only against the three worked examples in docs/designs/designs.md (KA-1..3, see
 * tests/test_oracle_known_answers.py) and by line-by-line restatement of the cited Go sources.
 *
 * Every function cites the reference file:line it follows (paths relative to the reference root).
 * Go `int`/`uint` are 64-bit on amd64: int64_t / uint64_t here, with C unsigned wrap-around
 * standing in for Go's.
 */
#ifndef GPUSHARE_ORACLE_H
#define GPUSHARE_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* v1.PodPhase */
enum {
    ORC_PHASE_PENDING = 0,
    ORC_PHASE_RUNNING = 1,
    ORC_PHASE_SUCCEEDED = 2,
    ORC_PHASE_FAILED = 3,
    ORC_PHASE_UNKNOWN = 4
};

/* Outcome of the gpushare predicate for one node (pkg/scheduler/gpushare-predicate.go:16-37). */
enum {
    ORC_FILTER_PASS = 0,
    ORC_FILTER_NOT_FOUND = 1,   /* node "<name>" not found                               */
    ORC_FILTER_NOT_SHARE = 2,   /* The node <name> is not for GPU share, need skip       */
    ORC_FILTER_INSUFFICIENT = 3 /* Insufficient GPU Memory in one device                 */
};

/* Scripted behaviour of the stand-in apiserver used by Allocate (pkg/cache/nodeinfo.go:139-206). */
enum {
    ORC_API_OK = 0,
    ORC_API_UPDATE_CONFLICT_THEN_OK = 1,      /* 1st Update -> optimistic-lock msg, GET ok, 2nd Update ok */
    ORC_API_UPDATE_CONFLICT_GET_FAILS = 2,    /* 1st Update -> optimistic-lock msg, GET fails             */
    ORC_API_UPDATE_CONFLICT_TWICE = 3,        /* both Updates fail                                        */
    ORC_API_UPDATE_OTHER_ERROR = 4,           /* Update fails with some other message                     */
    ORC_API_BIND_FAILS = 5                    /* Update ok, Bind fails                                    */
};

typedef struct orc_pod orc_pod;
typedef struct orc_cluster orc_cluster;

/* ---- strconv.Atoi (Go 1.10 stdlib; see .c for the restated algorithm) ---- */
/* *err: 0 = nil, 1 = ErrSyntax, 2 = ErrRange */
int64_t orc_atoi(const char* s, int* err);

/* ---- pod objects (the subset of v1.Pod the path reads) ---- */
orc_pod* orc_pod_new(const char* name, const char* ns, const char* uid);
orc_pod* orc_pod_clone(const orc_pod* p);
void orc_pod_free(orc_pod* p);
void orc_pod_set_node_name(orc_pod* p, const char* node_name);
void orc_pod_set_phase(orc_pod* p, int phase);
void orc_pod_set_deleting(orc_pod* p, int has_deletion_timestamp);
void orc_pod_set_annotation(orc_pod* p, const char* key, const char* value);
void orc_pod_clear_annotation(orc_pod* p, const char* key);
/* returns NULL when the key is absent */
const char* orc_pod_get_annotation(const orc_pod* p, const char* key);
/* appends a container; has_limit says whether limits["shared-gpu/gpu-mem"] exists */
void orc_pod_add_container(orc_pod* p, int has_limit, int64_t limit_value);
const char* orc_pod_node_name(const orc_pod* p);

/* ---- pkg/utils ---- */
int64_t orc_get_gpu_id_from_annotation(const orc_pod* p);            /* pod.go:45-60   */
uint64_t orc_get_gpu_memory_from_pod_annotation(const orc_pod* p);   /* pod.go:94-113  */
int64_t orc_get_gpu_memory_from_pod_resource(const orc_pod* p);      /* pod.go:146-155 */
int orc_is_complete_pod(const orc_pod* p);                           /* pod.go:28-37   */
int orc_is_gpu_sharing_pod(const orc_pod* p);                        /* pod.go:40-42   */
int orc_assigned_non_terminated_pod(const orc_pod* p);               /* pod.go:14-25   */

/* ---- cluster = node lister + pod lister + SchedulerCache + stand-in apiserver ---- */
orc_cluster* orc_cluster_new(void);
void orc_cluster_free(orc_cluster* c);

/* Node lister store.  has_mem/has_count say whether Status.Capacity holds the key. */
void orc_node_lister_set(orc_cluster* c, const char* name, int has_mem, int64_t mem, int has_count,
                         int64_t count);
void orc_node_lister_delete(orc_cluster* c, const char* name);
/* Pod lister store, keyed namespace/name; stores a clone. */
void orc_pod_lister_set(orc_cluster* c, const orc_pod* p);
void orc_pod_lister_delete(orc_cluster* c, const char* ns, const char* name);

/* SchedulerCache (pkg/cache/cache.go).  Return 0 on success; on error copy the message to err. */
int orc_build_cache(orc_cluster* c, char* err, size_t errcap);                        /* :49-74   */
int orc_add_or_update_pod(orc_cluster* c, const orc_pod* p, char* err, size_t errcap); /* :89-113  */
void orc_remove_pod(orc_cluster* c, const orc_pod* p);                                /* :116-127 */
int orc_known_pod(orc_cluster* c, const char* uid);                                   /* :81-87   */
/* GetNodeInfo (:130-162) exposed for its side effect (lazy NewNodeInfo / flip replacement). */
int orc_get_node_info(orc_cluster* c, const char* name, char* err, size_t errcap);

/* Filter: predicate closure + Predicate.Handler (pkg/scheduler/predicate.go:15-39). */
int orc_predicate(orc_cluster* c, const orc_pod* pod, const char* node_name, char* err,
                  size_t errcap); /* returns an ORC_FILTER_* code */
/* One Filter request; codes[i] is the outcome for node_names[i]. */
void orc_filter_codes(orc_cluster* c, const orc_pod* pod, const char* const* node_names, size_t n,
                      uint8_t* codes);
/* One Filter request rendered as the HTTP response body json.Marshal would produce
 * (pkg/routes/routes.go:85-97).  malloc'd; release with orc_free. */
char* orc_filter_json(orc_cluster* c, const orc_pod* pod, const char* const* node_names, size_t n);
/* Many Filter requests (one per pod), spread over `threads` OS threads the way net/http would
 * serve concurrent requests; codes is [n_pods][n] row-major.  Used for the timed CPU baseline. */
void orc_filter_many(orc_cluster* c, const orc_pod* const* pods, size_t n_pods,
                     const char* const* node_names, size_t n, int threads, uint8_t* codes);

/* Bind: Bind.Handler + gpushare-bind closure + NodeInfo.Allocate.
 * api_mode is an ORC_API_* script for this call.  Returns 0 when ExtenderBindingResult.Error == "",
 * else 1 with the message in err.  *dev_id receives allocateGPUID's choice (-1 = none). */
int orc_bind(orc_cluster* c, const char* pod_name, const char* pod_ns, const char* pod_uid,
             const char* node, int api_mode, int64_t* dev_id, char* err, size_t errcap);
/* allocateGPUID alone (pkg/cache/nodeinfo.go:209-252): no commit.  returns found (0/1). */
int orc_allocate_gpu_id(orc_cluster* c, const char* node, const orc_pod* pod, int64_t* dev_id);
/* The stand-in apiserver's copy of a pod (what Update wrote); NULL if never written. */
const orc_pod* orc_api_get_pod(orc_cluster* c, const char* ns, const char* name);
void orc_api_set_pod(orc_cluster* c, const orc_pod* p);
void orc_set_now_ns(orc_cluster* c, int64_t now_ns);
size_t orc_api_bind_count(orc_cluster* c);

/* Inspect (pkg/scheduler/inspect.go:33-71): per-device total/used for one cached node.
 * Returns the number of devices (gpuCount), or -1 with err when GetNodeInfo fails. */
int orc_inspect_node(orc_cluster* c, const char* name, uint64_t* total, uint64_t* used, int cap,
                     char* err, size_t errcap);

/* Event path: FilteringResourceEventHandler + controller handlers + syncPod
 * (pkg/gpushare/controller.go:77-100,174-332).  The caller updates the pod lister first, as the
 * informer does, then reports the event; orc_ctrl_drain runs syncPod for every queued key. */
void orc_informer_add(orc_cluster* c, const orc_pod* p);
void orc_informer_update(orc_cluster* c, const orc_pod* old_p, const orc_pod* new_p);
void orc_informer_delete(orc_cluster* c, const orc_pod* p);
int orc_ctrl_drain(orc_cluster* c, char* err, size_t errcap);
size_t orc_ctrl_queue_len(orc_cluster* c);

/* Bulk construction for the timed CPU baseline: lists nodes "node-%06d" (total/count < 0 = key
 * absent; listed[i] == 0 = not in the lister) and running pods "run-<j>" with their annotations
 * rendered as strings, then feeds every pod through AddOrUpdatePod (pods on unlisted nodes are
 * skipped, as syncPod would fail them).  Returns the number of pods added to the cache. */
size_t orc_bulk_load(orc_cluster* c, size_t n_nodes, const int64_t* node_total, const int64_t* node_count,
                     const uint8_t* node_listed, size_t n_pods, const int32_t* pod_node,
                     const int64_t* pod_gpu, const int64_t* pod_mem, const uint8_t* pod_phase);

void orc_free(void* p);

#ifdef __cplusplus
}
#endif
#endif
