// Host-side mirror of the reference's cache / predicate / bind / inspect interface (include/gpushare_ext.h).
// Pure host C++ over the C ABI of gpushare_b200.h: every decision comes from the CUDA kernels.
#include "gpushare_ext.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "gpushare_b200.h"
#include "kernels.h"

namespace {

// pkg/utils/const.go:8-12
const char* const kAnnIdx = "SHARED_GPU_MEM_IDX";
const char* const kAnnPod = "SHARED_GPU_MEM_POD";
const char* const kAnnDev = "SHARED_GPU_MEM_DEV";
const char* const kAnnAssigned = "SHARED_GPU_MEM_ASSIGNED";
const char* const kAnnAssumeTime = "SHARED_GPU_MEM_ASSUME_TIME";
// pkg/cache/nodeinfo.go:15
const char* const kOptimisticLockMsg =
    "the object has been modified; please apply your changes to the latest version and try again";

int fail(const std::string& msg) {
    gs::set_last_error(msg.c_str());
    return 1;
}

}  // namespace

struct gs_ext_pod {
    std::string name, ns, uid, node_name;
    int phase = GS_PHASE_PENDING;
    bool deleting = false;
    std::vector<std::pair<std::string, std::string>> annotations;  // insertion order, unique keys
    std::vector<std::pair<bool, int64_t>> containers;              // (has limit, limit)

    const std::string* annotation(const char* key) const {
        for (const auto& kv : annotations)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
    void set_annotation(const std::string& k, const std::string& v) {
        for (auto& kv : annotations)
            if (kv.first == k) {
                kv.second = v;
                return;
            }
        annotations.emplace_back(k, v);
    }
};

// strconv.Atoi (Go standard library, not part of the reference tree): optional sign, decimal digits only;
// syntax error -> 0; overflow is detected at the digit where it happens and clamps to the int64 range.
static int64_t go_atoi(const std::string& s, int* err) {
    *err = 0;
    size_t i = 0;
    bool neg = false;
    if (s.empty()) {
        *err = 1;
        return 0;
    }
    if (s[0] == '+' || s[0] == '-') {
        neg = s[0] == '-';
        i = 1;
        if (s.size() == 1) {
            *err = 1;
            return 0;
        }
    }
    unsigned long long n = 0;
    const unsigned long long cutoff = ~0ull / 10 + 1;
    bool range = false;
    for (; i < s.size(); ++i) {
        const unsigned char c = (unsigned char)s[i];
        if (c < '0' || c > '9') {
            *err = 1;
            return 0;
        }
        if (n >= cutoff) {
            range = true;
            break;
        }
        n *= 10;
        const unsigned long long n1 = n + (c - '0');
        if (n1 < n) {
            range = true;
            break;
        }
        n = n1;
    }
    if (range) n = ~0ull;
    const unsigned long long icut = 1ull << 63;
    if (!neg && n >= icut) {
        *err = 2;
        return INT64_MAX;
    }
    if (neg && n > icut) {
        *err = 2;
        return INT64_MIN;
    }
    if (range) *err = 2;
    return neg ? (int64_t)(0 - n) : (int64_t)n;
}

// pkg/utils/pod.go:45-60
static int64_t pod_gpu_id(const gs_ext_pod& p) {
    if (p.annotations.empty()) return -1;
    const std::string* v = p.annotation(kAnnIdx);
    if (!v) return -1;
    int err;
    const int64_t id = go_atoi(*v, &err);
    return err ? -1 : id;
}
// pkg/utils/pod.go:94-113 (the Atoi error is discarded there)
static uint64_t pod_annotation_mem(const gs_ext_pod& p) {
    if (p.annotations.empty()) return 0;
    const std::string* v = p.annotation(kAnnPod);
    if (!v) return 0;
    int err;
    int64_t s = go_atoi(*v, &err);
    if (s < 0) s = 0;
    return (uint64_t)s;
}
// pkg/utils/pod.go:146-155 (Go int addition wraps)
static int64_t pod_resource_mem(const gs_ext_pod& p) {
    uint64_t total = 0;
    for (const auto& c : p.containers)
        if (c.first) total += (uint64_t)c.second;
    return (int64_t)total;
}
static bool pod_counted(const gs_ext_pod& p) { return p.phase != GS_PHASE_SUCCEEDED && p.phase != GS_PHASE_FAILED; }

extern "C" {

gs_ext_pod* gs_ext_pod_new(const char* name, const char* ns, const char* uid) {
    gs_ext_pod* p = new gs_ext_pod();
    p->name = name ? name : "";
    p->ns = ns ? ns : "";
    p->uid = uid ? uid : "";
    return p;
}
void gs_ext_pod_free(gs_ext_pod* p) { delete p; }
void gs_ext_pod_set_node_name(gs_ext_pod* p, const char* node_name) { p->node_name = node_name ? node_name : ""; }
void gs_ext_pod_set_phase(gs_ext_pod* p, int phase) { p->phase = phase; }
void gs_ext_pod_set_deleting(gs_ext_pod* p, int d) { p->deleting = d != 0; }
void gs_ext_pod_set_annotation(gs_ext_pod* p, const char* key, const char* value) { p->set_annotation(key, value); }
void gs_ext_pod_add_container(gs_ext_pod* p, int has_limit, int64_t limit_value) {
    p->containers.emplace_back(has_limit != 0, limit_value);
}
int64_t gs_ext_atoi(const char* s, int* err) {
    int e;
    const int64_t v = go_atoi(s ? s : "", &e);
    if (err) *err = e;
    return v;
}
int64_t gs_ext_pod_gpu_id(const gs_ext_pod* p) { return pod_gpu_id(*p); }
uint64_t gs_ext_pod_annotation_mem(const gs_ext_pod* p) { return pod_annotation_mem(*p); }
int64_t gs_ext_pod_resource_mem(const gs_ext_pod* p) { return pod_resource_mem(*p); }
void gs_ext_free(void* p) { free(p); }

}  // extern "C"

#include "ext_state.inc"
