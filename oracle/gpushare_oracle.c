/*
 * gpushare_oracle.c -- CPU restatement of the gpushare-scheduler-extender hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see gpushare_oracle.h).  PARITY UNPINNED by reference tests: the
 * reference ships none; this file is pinned against docs/designs/designs.md KA-1..3 only.
 *
 * The cost structure deliberately follows the Go code: map-of-maps, `used` re-summed by parsing the
 * annotation string of every cached pod on every evaluation, three fresh maps per getAvailableGPUs,
 * node lookup by name under an exclusive lock.  Logging (beego log.Debug with %v of whole maps) is
 * compiled out, which favours the CPU baseline.
 */
#define _GNU_SOURCE
#include "gpushare_oracle.h"

#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* pkg/utils/const.go:3-13 */
#define RESOURCE_NAME "shared-gpu/gpu-mem"
#define COUNT_NAME "shared-gpu/gpu-count"
#define ENV_RESOURCE_INDEX "SHARED_GPU_MEM_IDX"
#define ENV_RESOURCE_BY_POD "SHARED_GPU_MEM_POD"
#define ENV_RESOURCE_BY_DEV "SHARED_GPU_MEM_DEV"
#define ENV_ASSIGNED_FLAG "SHARED_GPU_MEM_ASSIGNED"
#define ENV_ASSUME_TIME "SHARED_GPU_MEM_ASSUME_TIME"
/* pkg/cache/nodeinfo.go:15 */
#define OPTIMISTIC_LOCK_MSG \
    "the object has been modified; please apply your changes to the latest version and try again"

/* ------------------------------------------------------------------------------------------ */
/* small helpers                                                                              */
/* ------------------------------------------------------------------------------------------ */

static char* xstrdup(const char* s) {
    size_t n = strlen(s) + 1;
    char* r = (char*)malloc(n);
    memcpy(r, s, n);
    return r;
}

static void set_err(char* err, size_t cap, const char* msg) {
    if (err && cap) {
        snprintf(err, cap, "%s", msg);
    }
}

/* string -> void* hash map standing in for Go's map[string]T (open addressing, tombstones). */
typedef struct {
    char* key; /* NULL = empty, TOMB = deleted */
    void* val;
} sm_slot;
typedef struct {
    sm_slot* slots;
    size_t cap, len, used;
} strmap;
static char TOMB_[1];
#define TOMB ((char*)TOMB_)

static uint64_t sm_hash(const char* s) {
    uint64_t h = 1469598103934665603ull;
    for (; *s; ++s) {
        h ^= (unsigned char)*s;
        h *= 1099511628211ull;
    }
    return h;
}
static void sm_init(strmap* m) {
    m->cap = 8;
    m->len = m->used = 0;
    m->slots = (sm_slot*)calloc(m->cap, sizeof(sm_slot));
}
static void sm_put(strmap* m, const char* key, void* val);
static void sm_grow(strmap* m) {
    strmap n;
    n.cap = m->cap * 2;
    n.len = n.used = 0;
    n.slots = (sm_slot*)calloc(n.cap, sizeof(sm_slot));
    for (size_t i = 0; i < m->cap; ++i) {
        if (m->slots[i].key && m->slots[i].key != TOMB) {
            size_t j = sm_hash(m->slots[i].key) & (n.cap - 1);
            while (n.slots[j].key) j = (j + 1) & (n.cap - 1);
            n.slots[j] = m->slots[i];
            n.len++;
            n.used++;
        }
    }
    free(m->slots);
    *m = n;
}
static sm_slot* sm_find(const strmap* m, const char* key) {
    size_t j = sm_hash(key) & (m->cap - 1);
    while (m->slots[j].key) {
        if (m->slots[j].key != TOMB && strcmp(m->slots[j].key, key) == 0) return &m->slots[j];
        j = (j + 1) & (m->cap - 1);
    }
    return NULL;
}
static void* sm_get(const strmap* m, const char* key) {
    sm_slot* s = sm_find(m, key);
    return s ? s->val : NULL;
}
static int sm_has(const strmap* m, const char* key) { return sm_find(m, key) != NULL; }
/* returns the previous value (caller frees) or NULL */
static void* sm_swap(strmap* m, const char* key, void* val) {
    sm_slot* s = sm_find(m, key);
    if (s) {
        void* old = s->val;
        s->val = val;
        return old;
    }
    sm_put(m, key, val);
    return NULL;
}
static void sm_put(strmap* m, const char* key, void* val) {
    if ((m->used + 1) * 4 > m->cap * 3) sm_grow(m);
    size_t j = sm_hash(key) & (m->cap - 1);
    while (m->slots[j].key && m->slots[j].key != TOMB) j = (j + 1) & (m->cap - 1);
    if (!m->slots[j].key) m->used++;
    m->slots[j].key = xstrdup(key);
    m->slots[j].val = val;
    m->len++;
}
/* returns the removed value or NULL */
static void* sm_del(strmap* m, const char* key) {
    sm_slot* s = sm_find(m, key);
    if (!s) return NULL;
    void* v = s->val;
    free(s->key);
    s->key = TOMB;
    s->val = NULL;
    m->len--;
    return v;
}
static void sm_destroy(strmap* m, void (*free_val)(void*)) {
    for (size_t i = 0; i < m->cap; ++i) {
        if (m->slots[i].key && m->slots[i].key != TOMB) {
            free(m->slots[i].key);
            if (free_val && m->slots[i].val) free_val(m->slots[i].val);
        }
    }
    free(m->slots);
    m->slots = NULL;
    m->cap = m->len = m->used = 0;
}
#define SM_FOREACH(m, s) \
    for (sm_slot* s = (m)->slots; s < (m)->slots + (m)->cap; ++s) \
        if (s->key && s->key != TOMB)

/* ------------------------------------------------------------------------------------------ */
/* strconv.Atoi -- Go standard library (Go 1.10 per the reference's Dockerfile:1 and .travis.yml); */
/* NOT under /root/reference.  Published algorithm (strconv/atoi.go, ParseUint/ParseInt/Atoi):     */
/*   optional single '+'/'-', then one or more decimal digits, nothing else (no spaces, no '_');  */
/*   empty / sign only / any other byte -> (0, ErrSyntax);                                        */
/*   magnitude overflow is reported at the digit where it happens (before later bytes are looked  */
/*   at): positive -> (MaxInt64, ErrRange), negative -> (MinInt64, ErrRange).                     */
/* Call sites in the reference: pkg/utils/pod.go:51,81,98,133.                                    */
/* ------------------------------------------------------------------------------------------ */
int64_t orc_atoi(const char* s, int* err) {
    int e = 0;
    int64_t out = 0;
    const char* p = s;
    int neg = 0;
    if (*p == '\0') {
        e = 1;
        goto done;
    }
    if (*p == '+' || *p == '-') {
        neg = (*p == '-');
        ++p;
        if (*p == '\0') {
            e = 1;
            goto done;
        }
    }
    {
        /* ParseUint(s, 10, 64) */
        const uint64_t cutoff = UINT64_MAX / 10 + 1;
        uint64_t n = 0;
        for (; *p; ++p) {
            unsigned char d = (unsigned char)*p;
            if (d < '0' || d > '9') {
                e = 1;
                out = 0;
                goto done;
            }
            uint64_t v = (uint64_t)(d - '0');
            if (n >= cutoff) {
                n = UINT64_MAX;
                e = 2;
                break;
            }
            n *= 10;
            uint64_t n1 = n + v;
            if (n1 < n) {
                n = UINT64_MAX;
                e = 2;
                break;
            }
            n = n1;
        }
        /* ParseInt(s, 10, 0): clamp to the signed range */
        const uint64_t icut = (uint64_t)1 << 63;
        if (!neg && n >= icut) {
            out = INT64_MAX;
            e = 2;
        } else if (neg && n > icut) {
            out = INT64_MIN;
            e = 2;
        } else if (neg) {
            out = (int64_t)(0 - n);
        } else {
            out = (int64_t)n;
        }
    }
done:
    if (err) *err = e;
    return out;
}

/* ------------------------------------------------------------------------------------------ */
/* pods                                                                                       */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
    char* k;
    char* v;
} kv;
typedef struct {
    int has_limit;
    int64_t limit;
} ctr;
struct orc_pod {
    char *name, *ns, *uid, *node_name;
    int phase;
    int deleting;
    kv* ann;
    size_t n_ann;
    ctr* ctrs;
    size_t n_ctr;
};

orc_pod* orc_pod_new(const char* name, const char* ns, const char* uid) {
    orc_pod* p = (orc_pod*)calloc(1, sizeof(orc_pod));
    p->name = xstrdup(name);
    p->ns = xstrdup(ns);
    p->uid = xstrdup(uid);
    p->node_name = xstrdup("");
    p->phase = ORC_PHASE_PENDING;
    return p;
}
orc_pod* orc_pod_clone(const orc_pod* s) { /* stands in for pod.DeepCopy() */
    orc_pod* p = orc_pod_new(s->name, s->ns, s->uid);
    free(p->node_name);
    p->node_name = xstrdup(s->node_name);
    p->phase = s->phase;
    p->deleting = s->deleting;
    p->n_ann = s->n_ann;
    p->ann = (kv*)malloc(sizeof(kv) * (s->n_ann ? s->n_ann : 1));
    for (size_t i = 0; i < s->n_ann; ++i) {
        p->ann[i].k = xstrdup(s->ann[i].k);
        p->ann[i].v = xstrdup(s->ann[i].v);
    }
    p->n_ctr = s->n_ctr;
    p->ctrs = (ctr*)malloc(sizeof(ctr) * (s->n_ctr ? s->n_ctr : 1));
    memcpy(p->ctrs, s->ctrs, sizeof(ctr) * s->n_ctr);
    return p;
}
void orc_pod_free(orc_pod* p) {
    if (!p) return;
    free(p->name);
    free(p->ns);
    free(p->uid);
    free(p->node_name);
    for (size_t i = 0; i < p->n_ann; ++i) {
        free(p->ann[i].k);
        free(p->ann[i].v);
    }
    free(p->ann);
    free(p->ctrs);
    free(p);
}
static void pod_free_v(void* p) { orc_pod_free((orc_pod*)p); }
void orc_pod_set_node_name(orc_pod* p, const char* node_name) {
    free(p->node_name);
    p->node_name = xstrdup(node_name);
}
void orc_pod_set_phase(orc_pod* p, int phase) { p->phase = phase; }
void orc_pod_set_deleting(orc_pod* p, int d) { p->deleting = d; }
const char* orc_pod_node_name(const orc_pod* p) { return p->node_name; }
const char* orc_pod_get_annotation(const orc_pod* p, const char* key) {
    for (size_t i = 0; i < p->n_ann; ++i)
        if (strcmp(p->ann[i].k, key) == 0) return p->ann[i].v;
    return NULL;
}
void orc_pod_set_annotation(orc_pod* p, const char* key, const char* value) {
    for (size_t i = 0; i < p->n_ann; ++i) {
        if (strcmp(p->ann[i].k, key) == 0) {
            free(p->ann[i].v);
            p->ann[i].v = xstrdup(value);
            return;
        }
    }
    p->ann = (kv*)realloc(p->ann, sizeof(kv) * (p->n_ann + 1));
    p->ann[p->n_ann].k = xstrdup(key);
    p->ann[p->n_ann].v = xstrdup(value);
    p->n_ann++;
}
void orc_pod_clear_annotation(orc_pod* p, const char* key) {
    for (size_t i = 0; i < p->n_ann; ++i) {
        if (strcmp(p->ann[i].k, key) == 0) {
            free(p->ann[i].k);
            free(p->ann[i].v);
            p->ann[i] = p->ann[p->n_ann - 1];
            p->n_ann--;
            return;
        }
    }
}
void orc_pod_add_container(orc_pod* p, int has_limit, int64_t limit_value) {
    p->ctrs = (ctr*)realloc(p->ctrs, sizeof(ctr) * (p->n_ctr + 1));
    p->ctrs[p->n_ctr].has_limit = has_limit;
    p->ctrs[p->n_ctr].limit = limit_value;
    p->n_ctr++;
}

/* ------------------------------------------------------------------------------------------ */
/* pkg/utils/pod.go                                                                           */
/* ------------------------------------------------------------------------------------------ */

/* pod.go:14-25 AssignedNonTerminatedPod */
int orc_assigned_non_terminated_pod(const orc_pod* p) {
    if (p->deleting) return 0;
    if (p->node_name[0] == '\0') return 0;
    if (p->phase == ORC_PHASE_SUCCEEDED || p->phase == ORC_PHASE_FAILED) return 0;
    return 1;
}
/* pod.go:28-37 IsCompletePod */
int orc_is_complete_pod(const orc_pod* p) {
    if (p->deleting) return 1;
    if (p->phase == ORC_PHASE_SUCCEEDED || p->phase == ORC_PHASE_FAILED) return 1;
    return 0;
}
/* pod.go:45-60 GetGPUIDFromAnnotation: missing / any Atoi error -> -1 */
int64_t orc_get_gpu_id_from_annotation(const orc_pod* p) {
    int64_t id = -1;
    if (p->n_ann > 0) {
        const char* v = orc_pod_get_annotation(p, ENV_RESOURCE_INDEX);
        if (v) {
            int e;
            id = orc_atoi(v, &e);
            if (e) id = -1;
        }
    }
    return id;
}
/* pod.go:94-113 GetGPUMemoryFromPodAnnotation: the Atoi error is DISCARDED (`s, _ :=`), so a range
 * error yields MaxInt64 (or MinInt64 -> clamped to 0), a syntax error yields 0. */
uint64_t orc_get_gpu_memory_from_pod_annotation(const orc_pod* p) {
    uint64_t gpu_memory = 0;
    if (p->n_ann > 0) {
        const char* v = orc_pod_get_annotation(p, ENV_RESOURCE_BY_POD);
        if (v) {
            int e;
            int64_t s = orc_atoi(v, &e);
            if (s < 0) s = 0;
            gpu_memory += (uint64_t)s;
        }
    }
    return gpu_memory;
}
/* pod.go:146-155 GetGPUMemoryFromPodResource: sum over spec.containers (init containers ignored);
 * Quantity.Value() (vendor/k8s.io/apimachinery/pkg/api/resource/quantity.go:693-712) is taken as
 * the already-integral int64 the caller supplies. Go int addition wraps. */
int64_t orc_get_gpu_memory_from_pod_resource(const orc_pod* p) {
    uint64_t total = 0;
    for (size_t i = 0; i < p->n_ctr; ++i)
        if (p->ctrs[i].has_limit) total += (uint64_t)p->ctrs[i].limit;
    return (int64_t)total;
}
/* pod.go:40-42 IsGPUsharingPod */
int orc_is_gpu_sharing_pod(const orc_pod* p) { return orc_get_gpu_memory_from_pod_resource(p) > 0; }

/* pod.go:192-206 GetUpdatedPodAnnotationSpec */
static orc_pod* get_updated_pod_annotation_spec(const orc_pod* old_pod, int64_t dev_id,
                                                int64_t total_by_dev, int64_t now_ns) {
    orc_pod* np = orc_pod_clone(old_pod);
    char buf[32];
    snprintf(buf, sizeof buf, "%lld", (long long)dev_id);
    orc_pod_set_annotation(np, ENV_RESOURCE_INDEX, buf);
    snprintf(buf, sizeof buf, "%lld", (long long)total_by_dev);
    orc_pod_set_annotation(np, ENV_RESOURCE_BY_DEV, buf);
    snprintf(buf, sizeof buf, "%lld", (long long)orc_get_gpu_memory_from_pod_resource(np));
    orc_pod_set_annotation(np, ENV_RESOURCE_BY_POD, buf);
    orc_pod_set_annotation(np, ENV_ASSIGNED_FLAG, "false");
    snprintf(buf, sizeof buf, "%lld", (long long)now_ns);
    orc_pod_set_annotation(np, ENV_ASSUME_TIME, buf);
    return np;
}

/* ------------------------------------------------------------------------------------------ */
/* nodes: pkg/utils/node.go, pkg/cache/deviceinfo.go, pkg/cache/nodeinfo.go                    */
/* ------------------------------------------------------------------------------------------ */
typedef struct {
    char* name;
    int has_mem, has_count;
    int64_t mem, count;
} node_obj; /* the slice of v1.Node.Status.Capacity the path reads */

/* node.go:11-19 GetTotalGPUMemory */
static int64_t get_total_gpu_memory(const node_obj* n) { return n->has_mem ? n->mem : 0; }
/* node.go:22-30 GetGPUCountInNode */
static int64_t get_gpu_count_in_node(const node_obj* n) { return n->has_count ? n->count : 0; }
/* node.go:6-8 IsGPUSharingNode */
static int is_gpu_sharing_node(const node_obj* n) { return get_total_gpu_memory(n) > 0; }

/* deviceinfo.go:12-18 */
typedef struct {
    int64_t idx;
    strmap pod_map; /* uid -> orc_pod* (owned) */
    uint64_t total_gpu_mem;
    pthread_rwlock_t rwmu;
} device_info;

/* nodeinfo.go:19-26 */
typedef struct {
    char* name;
    node_obj node; /* the *v1.Node captured at creation: a frozen snapshot */
    device_info* devs;
    int64_t n_devs; /* len(n.devs): devs map has keys 0..n_devs-1 */
    int64_t gpu_count;
    int64_t gpu_total_memory;
    pthread_rwlock_t rwmu;
} node_info;

static void node_obj_copy(node_obj* d, const node_obj* s) {
    *d = *s;
    d->name = xstrdup(s->name);
}

/* nodeinfo.go:29-45 NewNodeInfo: one DeviceInfo per index below gpu-count, each with
 * totalGPUMem = uint(total / count) (Go int division truncates toward zero like C). */
static node_info* new_node_info(const node_obj* node) {
    node_info* n = (node_info*)calloc(1, sizeof(node_info));
    int64_t count = get_gpu_count_in_node(node);
    n->n_devs = count > 0 ? count : 0;
    n->devs = (device_info*)calloc((size_t)(n->n_devs ? n->n_devs : 1), sizeof(device_info));
    for (int64_t i = 0; i < count; ++i) {
        n->devs[i].idx = i;
        n->devs[i].total_gpu_mem = (uint64_t)(get_total_gpu_memory(node) / get_gpu_count_in_node(node));
        sm_init(&n->devs[i].pod_map);
        pthread_rwlock_init(&n->devs[i].rwmu, NULL);
    }
    n->name = xstrdup(node->name);
    node_obj_copy(&n->node, node);
    n->gpu_count = count;
    n->gpu_total_memory = get_total_gpu_memory(node);
    pthread_rwlock_init(&n->rwmu, NULL);
    return n;
}
static void node_info_free(void* v) {
    node_info* n = (node_info*)v;
    for (int64_t i = 0; i < n->n_devs; ++i) {
        sm_destroy(&n->devs[i].pod_map, pod_free_v);
        pthread_rwlock_destroy(&n->devs[i].rwmu);
    }
    free(n->devs);
    free(n->name);
    free(n->node.name);
    pthread_rwlock_destroy(&n->rwmu);
    free(n);
}

/* deviceinfo.go:41-54 GetUsedGPUMemory: re-sum the annotation of every pod in podMap, skipping
 * Succeeded / Failed; uint addition wraps. */
static uint64_t dev_get_used_gpu_memory(device_info* d) {
    uint64_t gpu_mem = 0;
    pthread_rwlock_rdlock(&d->rwmu);
    SM_FOREACH(&d->pod_map, s) {
        const orc_pod* pod = (const orc_pod*)s->val;
        if (pod->phase == ORC_PHASE_SUCCEEDED || pod->phase == ORC_PHASE_FAILED) continue;
        gpu_mem += orc_get_gpu_memory_from_pod_annotation(pod);
    }
    pthread_rwlock_unlock(&d->rwmu);
    return gpu_mem;
}
/* deviceinfo.go:56-67 addPod: podMap[pod.UID] = pod (replaces an entry with the same UID) */
static void dev_add_pod(device_info* d, orc_pod* pod_owned) {
    pthread_rwlock_wrlock(&d->rwmu);
    orc_pod* old = (orc_pod*)sm_swap(&d->pod_map, pod_owned->uid, pod_owned);
    pthread_rwlock_unlock(&d->rwmu);
    orc_pod_free(old);
}
/* deviceinfo.go:69-80 removePod: delete(podMap, pod.UID) */
static void dev_remove_pod(device_info* d, const orc_pod* pod) {
    pthread_rwlock_wrlock(&d->rwmu);
    orc_pod* old = (orc_pod*)sm_del(&d->pod_map, pod->uid);
    pthread_rwlock_unlock(&d->rwmu);
    orc_pod_free(old);
}

/* Go's map[int]uint for the three per-call maps of nodeinfo.go:254-284.  Keys are device indices
 * 0..n-1, so a presence-flagged array has the same observable behaviour; it is still allocated
 * fresh on every call, as the Go maps are. */
typedef struct {
    uint64_t* v;
    unsigned char* has;
    int64_t n;
} intmap;
static intmap im_new(int64_t n) {
    intmap m;
    m.n = n;
    m.v = (uint64_t*)calloc((size_t)(n ? n : 1), sizeof(uint64_t));
    m.has = (unsigned char*)calloc((size_t)(n ? n : 1), 1);
    return m;
}
static void im_free(intmap* m) {
    free(m->v);
    free(m->has);
}
static int64_t im_len(const intmap* m) {
    int64_t c = 0;
    for (int64_t i = 0; i < m->n; ++i) c += m->has[i];
    return c;
}
/* nodeinfo.go:277-284 getAllGPUs */
static intmap node_get_all_gpus(node_info* n) {
    intmap all = im_new(n->n_devs);
    for (int64_t i = 0; i < n->n_devs; ++i) {
        all.v[n->devs[i].idx] = n->devs[i].total_gpu_mem;
        all.has[n->devs[i].idx] = 1;
    }
    return all;
}
/* nodeinfo.go:267-274 getUsedGPUs */
static intmap node_get_used_gpus(node_info* n) {
    intmap used = im_new(n->n_devs);
    for (int64_t i = 0; i < n->n_devs; ++i) {
        used.v[n->devs[i].idx] = dev_get_used_gpu_memory(&n->devs[i]);
        used.has[n->devs[i].idx] = 1;
    }
    return used;
}
/* nodeinfo.go:254-264 getAvailableGPUs: total - used as Go uint (wraps mod 2^64) */
static intmap node_get_available_gpus(node_info* n) {
    intmap all = node_get_all_gpus(n);
    intmap used = node_get_used_gpus(n);
    intmap avail = im_new(n->n_devs);
    for (int64_t id = 0; id < all.n; ++id) {
        if (!all.has[id]) continue;
        if (used.has[id]) {
            avail.v[id] = all.v[id] - used.v[id];
            avail.has[id] = 1;
        }
    }
    im_free(&all);
    im_free(&used);
    return avail;
}

/* nodeinfo.go:113-137 Assume */
static int node_assume(node_info* n, const orc_pod* pod) {
    int allocatable = 0;
    pthread_rwlock_rdlock(&n->rwmu);
    intmap avail = node_get_available_gpus(n);
    uint64_t req = (uint64_t)orc_get_gpu_memory_from_pod_resource(pod);
    if (im_len(&avail) > 0) {
        for (int64_t dev_id = 0; dev_id < n->n_devs; ++dev_id) {
            if (avail.has[dev_id]) {
                if (avail.v[dev_id] >= req) {
                    allocatable = 1;
                    break;
                }
            }
        }
    }
    im_free(&avail);
    pthread_rwlock_unlock(&n->rwmu);
    return allocatable;
}

/* nodeinfo.go:209-252 allocateGPUID: req must be > 0; among fitting devices in ascending index
 * the one with strictly smaller available memory replaces the candidate (ties keep the lower). */
static int64_t node_allocate_gpu_id(node_info* n, const orc_pod* pod, int* found_out) {
    uint64_t req = 0;
    int found = 0;
    int64_t candidate_dev = -1;
    uint64_t candidate_mem = 0;
    intmap avail = node_get_available_gpus(n);
    req = (uint64_t)orc_get_gpu_memory_from_pod_resource(pod);
    if (req > 0) {
        if (im_len(&avail) > 0) {
            for (int64_t dev_id = 0; dev_id < n->n_devs; ++dev_id) {
                if (avail.has[dev_id]) {
                    if (avail.v[dev_id] >= req) {
                        if (candidate_dev == -1 || candidate_mem > avail.v[dev_id]) {
                            candidate_dev = dev_id;
                            candidate_mem = avail.v[dev_id];
                        }
                        found = 1;
                    }
                }
            }
        }
    }
    im_free(&avail);
    *found_out = found;
    return candidate_dev;
}

/* nodeinfo.go:89-110 addOrUpdatePod */
static int node_add_or_update_pod(node_info* n, orc_pod* pod_owned) {
    int added = 0;
    pthread_rwlock_wrlock(&n->rwmu);
    int64_t id = orc_get_gpu_id_from_annotation(pod_owned);
    if (id >= 0) {
        if (id < n->n_devs) { /* dev, found := n.devs[id] */
            dev_add_pod(&n->devs[id], pod_owned);
            added = 1;
        }
    }
    pthread_rwlock_unlock(&n->rwmu);
    return added;
}
/* nodeinfo.go:71-86 removePod */
static void node_remove_pod(node_info* n, const orc_pod* pod) {
    pthread_rwlock_wrlock(&n->rwmu);
    int64_t id = orc_get_gpu_id_from_annotation(pod);
    if (id >= 0) {
        if (id < n->n_devs) dev_remove_pod(&n->devs[id], pod);
    }
    pthread_rwlock_unlock(&n->rwmu);
}

/* ------------------------------------------------------------------------------------------ */
/* cluster: listers + pkg/cache/cache.go SchedulerCache + stand-in apiserver + controller state */
/* ------------------------------------------------------------------------------------------ */
struct orc_cluster {
    strmap node_lister; /* name -> node_obj*   */
    strmap pod_lister;  /* ns/name -> orc_pod* */
    /* SchedulerCache (cache.go:14-28) */
    strmap nodes;      /* name -> node_info*  */
    strmap known_pods; /* uid -> orc_pod*     */
    pthread_rwlock_t n_lock;
    /* stand-in apiserver */
    strmap api_pods; /* ns/name -> orc_pod* */
    size_t api_binds;
    int64_t now_ns;
    /* controller (controller.go:30-60): rate-limited workqueue reduced to an ordered set of keys,
     * plus removePodCache */
    char** queue;
    size_t q_len, q_cap;
    strmap remove_pod_cache; /* key -> orc_pod* */
};

static void node_obj_free(void* v) {
    node_obj* n = (node_obj*)v;
    free(n->name);
    free(n);
}

orc_cluster* orc_cluster_new(void) {
    orc_cluster* c = (orc_cluster*)calloc(1, sizeof(orc_cluster));
    sm_init(&c->node_lister);
    sm_init(&c->pod_lister);
    sm_init(&c->nodes);
    sm_init(&c->known_pods);
    sm_init(&c->api_pods);
    sm_init(&c->remove_pod_cache);
    pthread_rwlock_init(&c->n_lock, NULL);
    return c;
}
void orc_cluster_free(orc_cluster* c) {
    if (!c) return;
    sm_destroy(&c->node_lister, node_obj_free);
    sm_destroy(&c->pod_lister, pod_free_v);
    sm_destroy(&c->nodes, node_info_free);
    sm_destroy(&c->known_pods, pod_free_v);
    sm_destroy(&c->api_pods, pod_free_v);
    sm_destroy(&c->remove_pod_cache, pod_free_v);
    for (size_t i = 0; i < c->q_len; ++i) free(c->queue[i]);
    free(c->queue);
    pthread_rwlock_destroy(&c->n_lock);
    free(c);
}
void orc_free(void* p) { free(p); }
void orc_set_now_ns(orc_cluster* c, int64_t now_ns) { c->now_ns = now_ns; }
size_t orc_api_bind_count(orc_cluster* c) { return c->api_binds; }

void orc_node_lister_set(orc_cluster* c, const char* name, int has_mem, int64_t mem, int has_count,
                         int64_t count) {
    node_obj* n = (node_obj*)calloc(1, sizeof(node_obj));
    n->name = xstrdup(name);
    n->has_mem = has_mem;
    n->mem = mem;
    n->has_count = has_count;
    n->count = count;
    node_obj* old = (node_obj*)sm_swap(&c->node_lister, name, n);
    if (old) node_obj_free(old);
}
void orc_node_lister_delete(orc_cluster* c, const char* name) {
    node_obj* old = (node_obj*)sm_del(&c->node_lister, name);
    if (old) node_obj_free(old);
}
/* cache.MetaNamespaceKeyFunc: "<ns>/<name>", or "<name>" when the namespace is empty */
static char* pod_key(const char* ns, const char* name) {
    size_t a = strlen(ns), b = strlen(name);
    char* k = (char*)malloc(a + b + 2);
    if (a)
        snprintf(k, a + b + 2, "%s/%s", ns, name);
    else
        snprintf(k, a + b + 2, "%s", name);
    return k;
}
void orc_pod_lister_set(orc_cluster* c, const orc_pod* p) {
    char* k = pod_key(p->ns, p->name);
    orc_pod* old = (orc_pod*)sm_swap(&c->pod_lister, k, orc_pod_clone(p));
    orc_pod_free(old);
    free(k);
}
void orc_pod_lister_delete(orc_cluster* c, const char* ns, const char* name) {
    char* k = pod_key(ns, name);
    orc_pod_free((orc_pod*)sm_del(&c->pod_lister, k));
    free(k);
}
const orc_pod* orc_api_get_pod(orc_cluster* c, const char* ns, const char* name) {
    char* k = pod_key(ns, name);
    const orc_pod* p = (const orc_pod*)sm_get(&c->api_pods, k);
    free(k);
    return p;
}
void orc_api_set_pod(orc_cluster* c, const orc_pod* p) {
    char* k = pod_key(p->ns, p->name);
    orc_pod_free((orc_pod*)sm_swap(&c->api_pods, k, orc_pod_clone(p)));
    free(k);
}

/* strconv.Quote for the %q in the NotFound message
 * (vendor/k8s.io/apimachinery/pkg/api/errors/errors.go:101); ASCII printable names only need the
 * quote and backslash escapes. */
static void go_quote(const char* s, char* out, size_t cap) {
    size_t o = 0;
    if (cap < 3) {
        if (cap) out[0] = 0;
        return;
    }
    out[o++] = '"';
    for (; *s && o + 3 < cap; ++s) {
        if (*s == '"' || *s == '\\') out[o++] = '\\';
        out[o++] = *s;
    }
    out[o++] = '"';
    out[o] = 0;
}

/* cache.go:130-162 GetNodeInfo.  The node lister is consulted first; then, under the exclusive
 * nLock, the NodeInfo is created lazily, or REPLACED (dropping its podMaps) when the cached node had
 * total<=0 and the lister's has total>0, or cached count<=0 and the lister's count>0. */
static node_info* cache_get_node_info(orc_cluster* c, const char* name, char* err, size_t errcap) {
    const node_obj* node = (const node_obj*)sm_get(&c->node_lister, name);
    if (!node) {
        /* vendor/k8s.io/client-go/listers/core/v1/node.go:56-65 -> errors.NewNotFound */
        char q[512], msg[600];
        go_quote(name, q, sizeof q);
        snprintf(msg, sizeof msg, "node %s not found", q);
        set_err(err, errcap, msg);
        return NULL;
    }
    pthread_rwlock_wrlock(&c->n_lock);
    node_info* n = (node_info*)sm_get(&c->nodes, name);
    if (!n) {
        n = new_node_info(node);
        sm_put(&c->nodes, name, n);
    } else {
        if ((get_total_gpu_memory(&n->node) <= 0 && get_total_gpu_memory(node) > 0) ||
            (get_gpu_count_in_node(&n->node) <= 0 && get_gpu_count_in_node(node) > 0)) {
            node_info* fresh = new_node_info(node);
            node_info* old = (node_info*)sm_swap(&c->nodes, name, fresh);
            node_info_free(old);
            n = fresh;
        }
    }
    pthread_rwlock_unlock(&c->n_lock);
    return n;
}
int orc_get_node_info(orc_cluster* c, const char* name, char* err, size_t errcap) {
    return cache_get_node_info(c, name, err, errcap) ? 0 : -1;
}

/* cache.go:81-87 KnownPod */
int orc_known_pod(orc_cluster* c, const char* uid) {
    pthread_rwlock_rdlock(&c->n_lock);
    int f = sm_has(&c->known_pods, uid);
    pthread_rwlock_unlock(&c->n_lock);
    return f;
}
/* cache.go:170-174 rememberPod */
static void cache_remember_pod(orc_cluster* c, const orc_pod* pod) {
    pthread_rwlock_wrlock(&c->n_lock);
    orc_pod_free((orc_pod*)sm_swap(&c->known_pods, pod->uid, orc_pod_clone(pod)));
    pthread_rwlock_unlock(&c->n_lock);
}
/* cache.go:164-168 forgetPod */
static void cache_forget_pod(orc_cluster* c, const char* uid) {
    pthread_rwlock_wrlock(&c->n_lock);
    orc_pod_free((orc_pod*)sm_del(&c->known_pods, uid));
    pthread_rwlock_unlock(&c->n_lock);
}

/* cache.go:89-113 AddOrUpdatePod */
int orc_add_or_update_pod(orc_cluster* c, const orc_pod* pod, char* err, size_t errcap) {
    if (pod->node_name[0] == '\0') return 0;
    node_info* n = cache_get_node_info(c, pod->node_name, err, errcap);
    if (!n) return -1;
    orc_pod* pod_copy = orc_pod_clone(pod);
    if (node_add_or_update_pod(n, pod_copy)) {
        cache_remember_pod(c, pod);
    } else {
        orc_pod_free(pod_copy);
    }
    return 0;
}
/* cache.go:116-127 RemovePod */
void orc_remove_pod(orc_cluster* c, const orc_pod* pod) {
    node_info* n = cache_get_node_info(c, pod->node_name, NULL, 0);
    if (n) node_remove_pod(n, pod);
    cache_forget_pod(c, pod->uid);
}
/* cache.go:49-74 BuildCache: every listed pod with annotation memory > 0 and a node name. */
int orc_build_cache(orc_cluster* c, char* err, size_t errcap) {
    SM_FOREACH(&c->pod_lister, s) {
        const orc_pod* pod = (const orc_pod*)s->val;
        if (orc_get_gpu_memory_from_pod_annotation(pod) <= 0) continue;
        if (pod->node_name[0] == '\0') continue;
        if (orc_add_or_update_pod(c, pod, err, errcap) != 0) return -1;
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------ */
/* Filter: pkg/scheduler/gpushare-predicate.go:16-37 and pkg/scheduler/predicate.go:15-39       */
/* ------------------------------------------------------------------------------------------ */
int orc_predicate(orc_cluster* c, const orc_pod* pod, const char* node_name, char* err,
                  size_t errcap) {
    node_info* ni = cache_get_node_info(c, node_name, err, errcap);
    if (!ni) return ORC_FILTER_NOT_FOUND;
    if (!is_gpu_sharing_node(&ni->node)) {
        if (err && errcap)
            snprintf(err, errcap, "The node %s is not for GPU share, need skip", node_name);
        return ORC_FILTER_NOT_SHARE;
    }
    if (!node_assume(ni, pod)) {
        set_err(err, errcap, "Insufficient GPU Memory in one device");
        return ORC_FILTER_INSUFFICIENT;
    }
    return ORC_FILTER_PASS;
}

void orc_filter_codes(orc_cluster* c, const orc_pod* pod, const char* const* node_names, size_t n,
                      uint8_t* codes) {
    for (size_t i = 0; i < n; ++i) codes[i] = (uint8_t)orc_predicate(c, pod, node_names[i], NULL, 0);
}

typedef struct {
    orc_cluster* c;
    const orc_pod* const* pods;
    size_t p0, p1;
    const char* const* names;
    size_t n;
    uint8_t* codes;
} fm_job;
static void* fm_worker(void* v) {
    fm_job* j = (fm_job*)v;
    for (size_t p = j->p0; p < j->p1; ++p)
        orc_filter_codes(j->c, j->pods[p], j->names, j->n, j->codes + p * j->n);
    return NULL;
}
void orc_filter_many(orc_cluster* c, const orc_pod* const* pods, size_t n_pods,
                     const char* const* node_names, size_t n, int threads, uint8_t* codes) {
    if (threads < 1) threads = 1;
    if ((size_t)threads > n_pods) threads = n_pods ? (int)n_pods : 1;
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
    fm_job* jobs = (fm_job*)malloc(sizeof(fm_job) * (size_t)threads);
    for (int t = 0; t < threads; ++t) {
        jobs[t].c = c;
        jobs[t].pods = pods;
        jobs[t].p0 = n_pods * (size_t)t / (size_t)threads;
        jobs[t].p1 = n_pods * (size_t)(t + 1) / (size_t)threads;
        jobs[t].names = node_names;
        jobs[t].n = n;
        jobs[t].codes = codes;
        if (t > 0) pthread_create(&th[t], NULL, fm_worker, &jobs[t]);
    }
    fm_worker(&jobs[0]);
    for (int t = 1; t < threads; ++t) pthread_join(th[t], NULL);
    free(th);
    free(jobs);
}

/* growing string buffer for the JSON envelope */
typedef struct {
    char* p;
    size_t len, cap;
} sbuf;
static void sb_put(sbuf* b, const char* s, size_t n) {
    if (b->len + n + 1 > b->cap) {
        while (b->len + n + 1 > b->cap) b->cap = b->cap ? b->cap * 2 : 256;
        b->p = (char*)realloc(b->p, b->cap);
    }
    memcpy(b->p + b->len, s, n);
    b->len += n;
    b->p[b->len] = 0;
}
static void sb_puts(sbuf* b, const char* s) { sb_put(b, s, strlen(s)); }
/* encoding/json string encoding (Go stdlib, escapeHTML on): ", \, control bytes, <, >, & */
static void sb_json_string(sbuf* b, const char* s) {
    sb_puts(b, "\"");
    for (; *s; ++s) {
        unsigned char ch = (unsigned char)*s;
        char tmp[8];
        if (ch == '"' || ch == '\\') {
            tmp[0] = '\\';
            tmp[1] = (char)ch;
            sb_put(b, tmp, 2);
        } else if (ch == '\n') {
            sb_puts(b, "\\n");
        } else if (ch == '\r') {
            sb_puts(b, "\\r");
        } else if (ch == '\t') {
            sb_puts(b, "\\t");
        } else if (ch < 0x20 || ch == '<' || ch == '>' || ch == '&') {
            snprintf(tmp, sizeof tmp, "\\u%04x", ch);
            sb_puts(b, tmp);
        } else {
            sb_put(b, (const char*)&ch, 1);
        }
    }
    sb_puts(b, "\"");
}
static int cmp_kv(const void* a, const void* b) {
    return strcmp(((const kv*)a)->k, ((const kv*)b)->k);
}

/* Predicate.Handler (predicate.go:15-39) + json.Marshal of ExtenderFilterResult
 * (routes.go:85-97; vendor/k8s.io/kubernetes/pkg/scheduler/api/types.go:269-284): fields in struct
 * order with Go names, pass list in input order (duplicates kept), FailedNodes as a map (one entry
 * per distinct name, last error wins, keys sorted by encoding/json). */
char* orc_filter_json(orc_cluster* c, const orc_pod* pod, const char* const* node_names, size_t n) {
    sbuf b = {0};
    strmap failed; /* name -> char* message */
    sm_init(&failed);
    sb_puts(&b, "{\"Nodes\":null,\"NodeNames\":[");
    int first = 1;
    char err[1024];
    for (size_t i = 0; i < n; ++i) {
        int code = orc_predicate(c, pod, node_names[i], err, sizeof err);
        if (code != ORC_FILTER_PASS) {
            free(sm_swap(&failed, node_names[i], xstrdup(err)));
        } else {
            if (!first) sb_puts(&b, ",");
            first = 0;
            sb_json_string(&b, node_names[i]);
        }
    }
    sb_puts(&b, "],\"FailedNodes\":{");
    kv* items = (kv*)malloc(sizeof(kv) * (failed.len ? failed.len : 1));
    size_t k = 0;
    SM_FOREACH(&failed, s) {
        items[k].k = s->key;
        items[k].v = (char*)s->val;
        k++;
    }
    qsort(items, k, sizeof(kv), cmp_kv);
    for (size_t i = 0; i < k; ++i) {
        if (i) sb_puts(&b, ",");
        sb_json_string(&b, items[i].k);
        sb_puts(&b, ":");
        sb_json_string(&b, items[i].v);
    }
    free(items);
    sb_puts(&b, "},\"Error\":\"\"}");
    sm_destroy(&failed, free);
    return b.p;
}

/* ------------------------------------------------------------------------------------------ */
/* Bind: pkg/scheduler/bind.go:17-26, gpushare-bind.go:21-65, pkg/cache/nodeinfo.go:139-206     */
/* ------------------------------------------------------------------------------------------ */
int orc_allocate_gpu_id(orc_cluster* c, const char* node, const orc_pod* pod, int64_t* dev_id) {
    node_info* ni = cache_get_node_info(c, node, NULL, 0);
    int found = 0;
    *dev_id = -1;
    if (!ni) return 0;
    pthread_rwlock_wrlock(&ni->rwmu);
    *dev_id = node_allocate_gpu_id(ni, pod, &found);
    pthread_rwlock_unlock(&ni->rwmu);
    return found;
}

/* stand-in for clientset.CoreV1().Pods(ns).Update: stores the object */
static void api_update(orc_cluster* c, const orc_pod* p) { orc_api_set_pod(c, p); }

/* nodeinfo.go:139-206 Allocate */
static int node_allocate(orc_cluster* c, node_info* n, const orc_pod* pod_in, int api_mode,
                         int64_t* dev_out, char* err, size_t errcap) {
    int failed = 0;
    orc_pod* new_pod = NULL;
    orc_pod* refetched = NULL;
    const orc_pod* pod = pod_in;
    pthread_rwlock_wrlock(&n->rwmu);
    int found = 0;
    int64_t dev_id = node_allocate_gpu_id(n, pod, &found);
    *dev_out = found ? dev_id : -1;
    if (found) {
        /* n.GetTotalGPUMemory()/n.GetGPUCount() on the node-level ints */
        int64_t by_dev = n->gpu_total_memory / n->gpu_count;
        new_pod = get_updated_pod_annotation_spec(pod, dev_id, by_dev, c->now_ns);
        int update_failed = (api_mode == ORC_API_UPDATE_CONFLICT_THEN_OK ||
                             api_mode == ORC_API_UPDATE_CONFLICT_GET_FAILS ||
                             api_mode == ORC_API_UPDATE_CONFLICT_TWICE ||
                             api_mode == ORC_API_UPDATE_OTHER_ERROR);
        if (!update_failed) {
            api_update(c, new_pod);
        } else if (api_mode == ORC_API_UPDATE_OTHER_ERROR) {
            set_err(err, errcap, "stand-in apiserver: update failed");
            failed = 1;
            goto out;
        } else {
            /* err.Error() == OptimisticLockErrorMsg: re-GET, re-annotate with the SAME devId */
            if (api_mode == ORC_API_UPDATE_CONFLICT_GET_FAILS) {
                set_err(err, errcap, "stand-in apiserver: get failed");
                failed = 1;
                goto out;
            }
            const orc_pod* got = orc_api_get_pod(c, pod->ns, pod->name);
            refetched = orc_pod_clone(got ? got : pod);
            pod = refetched;
            orc_pod_free(new_pod);
            new_pod = get_updated_pod_annotation_spec(pod, dev_id, by_dev, c->now_ns);
            if (api_mode == ORC_API_UPDATE_CONFLICT_TWICE) {
                set_err(err, errcap, OPTIMISTIC_LOCK_MSG);
                failed = 1;
                goto out;
            }
            api_update(c, new_pod);
        }
    } else {
        /* nodeinfo.go:170 -- pod.Spec.NodeName, not the target node */
        if (err && errcap)
            snprintf(err, errcap, "The node %s can't place the pod %s in ns %s", pod->node_name,
                     pod->name, pod->ns);
        failed = 1;
    }
    /* 2. Bind */
    if (!failed) {
        if (api_mode == ORC_API_BIND_FAILS) {
            set_err(err, errcap, "stand-in apiserver: bind failed");
            failed = 1;
            goto out;
        }
        c->api_binds++;
    }
    /* 3. commit: dev.addPod(newPod) */
    if (!failed) {
        if (dev_id >= 0 && dev_id < n->n_devs) {
            dev_add_pod(&n->devs[dev_id], new_pod);
            new_pod = NULL;
        }
    }
out:
    pthread_rwlock_unlock(&n->rwmu);
    orc_pod_free(new_pod);
    orc_pod_free(refetched);
    return failed;
}

/* gpushare-bind.go:44-65 getPod */
static orc_pod* bind_get_pod(orc_cluster* c, const char* name, const char* ns, const char* uid,
                             char* err, size_t errcap) {
    char* k = pod_key(ns, name);
    const orc_pod* pod = (const orc_pod*)sm_get(&c->pod_lister, k);
    if (!pod) {
        char msg[600];
        snprintf(msg, sizeof msg, "pod \"%s\" not found", name);
        set_err(err, errcap, msg);
        free(k);
        return NULL;
    }
    if (strcmp(pod->uid, uid) != 0) {
        pod = (const orc_pod*)sm_get(&c->api_pods, k);
        if (!pod) {
            char msg[600];
            snprintf(msg, sizeof msg, "pods \"%s\" not found", name);
            set_err(err, errcap, msg);
            free(k);
            return NULL;
        }
        if (strcmp(pod->uid, uid) != 0) {
            if (err && errcap)
                snprintf(err, errcap,
                         "The pod %s in ns %s's uid is %s, and it's not equal with expected %s", name,
                         ns, pod->uid, uid);
            free(k);
            return NULL;
        }
    }
    free(k);
    return orc_pod_clone(pod);
}

int orc_bind(orc_cluster* c, const char* pod_name, const char* pod_ns, const char* pod_uid,
             const char* node, int api_mode, int64_t* dev_id, char* err, size_t errcap) {
    int64_t dev = -1;
    if (dev_id) *dev_id = -1;
    orc_pod* pod = bind_get_pod(c, pod_name, pod_ns, pod_uid, err, errcap);
    if (!pod) return 1;
    node_info* ni = cache_get_node_info(c, node, err, errcap);
    if (!ni) {
        orc_pod_free(pod);
        return 1;
    }
    int failed = node_allocate(c, ni, pod, api_mode, &dev, err, errcap);
    if (dev_id) *dev_id = dev;
    orc_pod_free(pod);
    return failed;
}

/* ------------------------------------------------------------------------------------------ */
/* Inspect: pkg/scheduler/inspect.go:33-71 (per-device totals only)                            */
/* ------------------------------------------------------------------------------------------ */
int orc_inspect_node(orc_cluster* c, const char* name, uint64_t* total, uint64_t* used, int cap,
                     char* err, size_t errcap) {
    node_info* ni = cache_get_node_info(c, name, err, errcap);
    if (!ni) return -1;
    for (int64_t i = 0; i < ni->n_devs && i < cap; ++i) {
        total[i] = ni->devs[i].total_gpu_mem;
        used[i] = dev_get_used_gpu_memory(&ni->devs[i]);
    }
    return (int)ni->gpu_count;
}

/* ------------------------------------------------------------------------------------------ */
/* Event path: pkg/gpushare/controller.go                                                      */
/* ------------------------------------------------------------------------------------------ */
static void queue_add(orc_cluster* c, const char* key) {
    /* workqueue.Add: a key already waiting is not queued twice */
    for (size_t i = 0; i < c->q_len; ++i)
        if (strcmp(c->queue[i], key) == 0) return;
    if (c->q_len == c->q_cap) {
        c->q_cap = c->q_cap ? c->q_cap * 2 : 16;
        c->queue = (char**)realloc(c->queue, sizeof(char*) * c->q_cap);
    }
    c->queue[c->q_len++] = xstrdup(key);
}
size_t orc_ctrl_queue_len(orc_cluster* c) { return c->q_len; }

/* controller.go:233-255 addPodToCache: always enqueue */
static void ctrl_add_pod_to_cache(orc_cluster* c, const orc_pod* pod) {
    char* k = pod_key(pod->ns, pod->name);
    queue_add(c, k);
    free(k);
}
/* controller.go:257-305 updatePodInCache */
static void ctrl_update_pod_in_cache(orc_cluster* c, const orc_pod* old_pod, const orc_pod* new_pod) {
    int need_update = 0;
    const char* pod_uid = old_pod->uid;
    if (orc_known_pod(c, pod_uid) && orc_is_complete_pod(new_pod)) need_update = 1;
    if (!orc_known_pod(c, pod_uid) && orc_get_gpu_id_from_annotation(new_pod) >= 0) need_update = 1;
    if (need_update) {
        char* k = pod_key(new_pod->ns, new_pod->name);
        queue_add(c, k);
        free(k);
    }
}
/* controller.go:307-332 deletePodFromCache: enqueue and stash the last-seen object */
static void ctrl_delete_pod_from_cache(orc_cluster* c, const orc_pod* pod) {
    char* k = pod_key(pod->ns, pod->name);
    queue_add(c, k);
    orc_pod_free((orc_pod*)sm_swap(&c->remove_pod_cache, k, orc_pod_clone(pod)));
    free(k);
}
/* FilteringResourceEventHandler (vendor/k8s.io/client-go/tools/cache/controller.go:215-252) with
 * FilterFunc = IsGPUsharingPod (controller.go:77-100) */
void orc_informer_add(orc_cluster* c, const orc_pod* p) {
    if (!orc_is_gpu_sharing_pod(p)) return;
    ctrl_add_pod_to_cache(c, p);
}
void orc_informer_update(orc_cluster* c, const orc_pod* old_p, const orc_pod* new_p) {
    int newer = orc_is_gpu_sharing_pod(new_p);
    int older = orc_is_gpu_sharing_pod(old_p);
    if (newer && older)
        ctrl_update_pod_in_cache(c, old_p, new_p);
    else if (newer && !older)
        ctrl_add_pod_to_cache(c, new_p);
    else if (!newer && older)
        ctrl_delete_pod_from_cache(c, old_p);
}
void orc_informer_delete(orc_cluster* c, const orc_pod* p) {
    if (!orc_is_gpu_sharing_pod(p)) return;
    ctrl_delete_pod_from_cache(c, p);
}
/* controller.go:174-205 syncPod */
static int ctrl_sync_pod(orc_cluster* c, const char* key, char* err, size_t errcap) {
    const orc_pod* pod = (const orc_pod*)sm_get(&c->pod_lister, key);
    if (!pod) {
        orc_pod* stashed = (orc_pod*)sm_del(&c->remove_pod_cache, key);
        if (stashed) {
            orc_remove_pod(c, stashed);
            orc_pod_free(stashed);
        }
    } else if (orc_is_complete_pod(pod)) {
        orc_remove_pod(c, pod);
    } else {
        if (orc_add_or_update_pod(c, pod, err, errcap) != 0) return -1;
    }
    return 0;
}
/* Drains the queue in order.  (The live reference handles ~1 key/s because processNextWorkItem
 * returns false on success, controller.go:209-224; that pacing is not a semantic and is not
 * restated.)  A key whose sync fails is dropped here and the error returned. */
int orc_ctrl_drain(orc_cluster* c, char* err, size_t errcap) {
    int rc = 0;
    while (c->q_len) {
        char* key = c->queue[0];
        memmove(c->queue, c->queue + 1, sizeof(char*) * (c->q_len - 1));
        c->q_len--;
        if (ctrl_sync_pod(c, key, err, errcap) != 0) rc = -1;
        free(key);
    }
    return rc;
}

/* ------------------------------------------------------------------------------------------ */
/* bulk construction (test / baseline plumbing, no reference counterpart)                       */
/* ------------------------------------------------------------------------------------------ */
size_t orc_bulk_load(orc_cluster* c, size_t n_nodes, const int64_t* node_total, const int64_t* node_count,
                     const uint8_t* node_listed, size_t n_pods, const int32_t* pod_node,
                     const int64_t* pod_gpu, const int64_t* pod_mem, const uint8_t* pod_phase) {
    char name[64], buf[32];
    for (size_t i = 0; i < n_nodes; ++i) {
        if (!node_listed[i]) continue;
        snprintf(name, sizeof name, "node-%06zu", i);
        orc_node_lister_set(c, name, node_total[i] >= 0, node_total[i] >= 0 ? node_total[i] : 0,
                            node_count[i] >= 0, node_count[i] >= 0 ? node_count[i] : 0);
    }
    size_t added = 0;
    for (size_t j = 0; j < n_pods; ++j) {
        char pname[64], uid[64];
        snprintf(pname, sizeof pname, "run-%zu", j);
        snprintf(uid, sizeof uid, "uid-run-%zu", j);
        snprintf(name, sizeof name, "node-%06zu", (size_t)pod_node[j]);
        orc_pod* p = orc_pod_new(pname, "default", uid);
        orc_pod_set_node_name(p, name);
        orc_pod_set_phase(p, pod_phase[j]);
        snprintf(buf, sizeof buf, "%lld", (long long)pod_gpu[j]);
        orc_pod_set_annotation(p, ENV_RESOURCE_INDEX, buf);
        snprintf(buf, sizeof buf, "%lld", (long long)pod_mem[j]);
        orc_pod_set_annotation(p, ENV_RESOURCE_BY_POD, buf);
        orc_pod_add_container(p, 1, pod_mem[j]);
        if (orc_get_gpu_memory_from_pod_annotation(p) > 0 && orc_add_or_update_pod(c, p, NULL, 0) == 0) added++;
        orc_pod_free(p);
    }
    return added;
}
