"""ctypes front-end of the CPU oracle (oracle/gpushare_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs.  The product package never imports this module.

Parity status: unpinned by reference tests (the reference has none); pinned against the worked
examples of docs/designs/designs.md only (tests/test_oracle_known_answers.py).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Iterable, Sequence

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libgpushare_oracle.so")

PENDING, RUNNING, SUCCEEDED, FAILED, UNKNOWN = range(5)
PASS, NOT_FOUND, NOT_SHARE, INSUFFICIENT = range(4)
API_OK, API_CONFLICT_THEN_OK, API_CONFLICT_GET_FAILS, API_CONFLICT_TWICE, API_OTHER_ERROR, API_BIND_FAILS = range(6)

IDX = "SHARED_GPU_MEM_IDX"
POD = "SHARED_GPU_MEM_POD"
DEV = "SHARED_GPU_MEM_DEV"
ASSIGNED = "SHARED_GPU_MEM_ASSIGNED"
ASSUME_TIME = "SHARED_GPU_MEM_ASSUME_TIME"


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (gcc only)."""
    src = os.path.join(_HERE, "gpushare_oracle.c")
    hdr = os.path.join(_HERE, "gpushare_oracle.h")
    stale = (not os.path.exists(_SO)) or any(
        os.path.exists(p) and os.path.getmtime(p) > os.path.getmtime(_SO) for p in (src, hdr)
    )
    if force or stale:
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(_SO)
    vp, cp, i64, u64, sz = C.c_void_p, C.c_char_p, C.c_int64, C.c_uint64, C.c_size_t
    sig = {
        "orc_atoi": (i64, [cp, C.POINTER(C.c_int)]),
        "orc_pod_new": (vp, [cp, cp, cp]),
        "orc_pod_clone": (vp, [vp]),
        "orc_pod_free": (None, [vp]),
        "orc_pod_set_node_name": (None, [vp, cp]),
        "orc_pod_set_phase": (None, [vp, C.c_int]),
        "orc_pod_set_deleting": (None, [vp, C.c_int]),
        "orc_pod_set_annotation": (None, [vp, cp, cp]),
        "orc_pod_clear_annotation": (None, [vp, cp]),
        "orc_pod_get_annotation": (cp, [vp, cp]),
        "orc_pod_add_container": (None, [vp, C.c_int, i64]),
        "orc_pod_node_name": (cp, [vp]),
        "orc_get_gpu_id_from_annotation": (i64, [vp]),
        "orc_get_gpu_memory_from_pod_annotation": (u64, [vp]),
        "orc_get_gpu_memory_from_pod_resource": (i64, [vp]),
        "orc_is_complete_pod": (C.c_int, [vp]),
        "orc_is_gpu_sharing_pod": (C.c_int, [vp]),
        "orc_assigned_non_terminated_pod": (C.c_int, [vp]),
        "orc_cluster_new": (vp, []),
        "orc_cluster_free": (None, [vp]),
        "orc_node_lister_set": (None, [vp, cp, C.c_int, i64, C.c_int, i64]),
        "orc_node_lister_delete": (None, [vp, cp]),
        "orc_pod_lister_set": (None, [vp, vp]),
        "orc_pod_lister_delete": (None, [vp, cp, cp]),
        "orc_build_cache": (C.c_int, [vp, cp, sz]),
        "orc_add_or_update_pod": (C.c_int, [vp, vp, cp, sz]),
        "orc_remove_pod": (None, [vp, vp]),
        "orc_known_pod": (C.c_int, [vp, cp]),
        "orc_get_node_info": (C.c_int, [vp, cp, cp, sz]),
        "orc_predicate": (C.c_int, [vp, vp, cp, cp, sz]),
        "orc_filter_codes": (None, [vp, vp, C.POINTER(cp), sz, C.POINTER(C.c_uint8)]),
        "orc_filter_json": (vp, [vp, vp, C.POINTER(cp), sz]),
        "orc_filter_many": (None, [vp, C.POINTER(vp), sz, C.POINTER(cp), sz, C.c_int, C.POINTER(C.c_uint8)]),
        "orc_bind": (C.c_int, [vp, cp, cp, cp, cp, C.c_int, C.POINTER(i64), cp, sz]),
        "orc_allocate_gpu_id": (C.c_int, [vp, cp, vp, C.POINTER(i64)]),
        "orc_api_get_pod": (vp, [vp, cp, cp]),
        "orc_api_set_pod": (None, [vp, vp]),
        "orc_set_now_ns": (None, [vp, i64]),
        "orc_api_bind_count": (sz, [vp]),
        "orc_inspect_node": (C.c_int, [vp, cp, C.POINTER(u64), C.POINTER(u64), C.c_int, cp, sz]),
        "orc_informer_add": (None, [vp, vp]),
        "orc_informer_update": (None, [vp, vp, vp]),
        "orc_informer_delete": (None, [vp, vp]),
        "orc_ctrl_drain": (C.c_int, [vp, cp, sz]),
        "orc_ctrl_queue_len": (sz, [vp]),
        "orc_bulk_load": (sz, [vp, sz, vp, vp, vp, sz, vp, vp, vp, vp]),
        "orc_free": (None, [vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _b(s: str) -> bytes:
    return s.encode("utf-8")


def atoi(s: str):
    """strconv.Atoi restatement -> (value, err) with err in {0: nil, 1: syntax, 2: range}."""
    e = C.c_int(0)
    v = lib().orc_atoi(_b(s), C.byref(e))
    return v, e.value


class Pod:
    """The slice of v1.Pod the hot path reads."""

    def __init__(self, name: str, ns: str = "default", uid: str | None = None, *, node: str = "",
                 phase: int = PENDING, deleting: bool = False, annotations: dict | None = None,
                 limits: Iterable | None = None, _h=None):
        L = lib()
        if _h is not None:
            self.h = _h
            return
        self.h = L.orc_pod_new(_b(name), _b(ns), _b(uid if uid is not None else f"uid-{ns}-{name}"))
        L.orc_pod_set_node_name(self.h, _b(node))
        L.orc_pod_set_phase(self.h, phase)
        L.orc_pod_set_deleting(self.h, 1 if deleting else 0)
        for k, v in (annotations or {}).items():
            L.orc_pod_set_annotation(self.h, _b(k), _b(str(v)))
        for lim in (limits if limits is not None else []):
            if lim is None:
                L.orc_pod_add_container(self.h, 0, 0)
            else:
                L.orc_pod_add_container(self.h, 1, int(lim))

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().orc_pod_free(self.h)
                self.h = None
        except Exception:
            pass

    def clone(self) -> "Pod":
        return Pod("", _h=lib().orc_pod_clone(self.h))

    def set_annotation(self, k, v):
        lib().orc_pod_set_annotation(self.h, _b(k), _b(str(v)))
        return self

    def clear_annotation(self, k):
        lib().orc_pod_clear_annotation(self.h, _b(k))
        return self

    def annotation(self, k):
        r = lib().orc_pod_get_annotation(self.h, _b(k))
        return None if r is None else r.decode()

    def set_node(self, node):
        lib().orc_pod_set_node_name(self.h, _b(node))
        return self

    def set_phase(self, phase):
        lib().orc_pod_set_phase(self.h, phase)
        return self

    def set_deleting(self, d=True):
        lib().orc_pod_set_deleting(self.h, 1 if d else 0)
        return self

    @property
    def gpu_id(self):
        return lib().orc_get_gpu_id_from_annotation(self.h)

    @property
    def annotation_mem(self):
        return lib().orc_get_gpu_memory_from_pod_annotation(self.h)

    @property
    def resource_mem(self):
        return lib().orc_get_gpu_memory_from_pod_resource(self.h)

    @property
    def complete(self):
        return bool(lib().orc_is_complete_pod(self.h))

    @property
    def sharing(self):
        return bool(lib().orc_is_gpu_sharing_pod(self.h))


def _names(names: Sequence[str]):
    arr = (C.c_char_p * len(names))(*[_b(n) for n in names])
    return arr


class Cluster:
    """Node lister + pod lister + SchedulerCache + stand-in apiserver + controller queue."""

    def __init__(self):
        self.h = lib().orc_cluster_new()
        self._err = C.create_string_buffer(2048)

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().orc_cluster_free(self.h)
                self.h = None
        except Exception:
            pass

    # listers
    def set_node(self, name, mem=None, count=None):
        lib().orc_node_lister_set(self.h, _b(name), mem is not None, int(mem or 0), count is not None, int(count or 0))

    def delete_node(self, name):
        lib().orc_node_lister_delete(self.h, _b(name))

    def list_pod(self, pod: Pod):
        lib().orc_pod_lister_set(self.h, pod.h)

    def unlist_pod(self, ns, name):
        lib().orc_pod_lister_delete(self.h, _b(ns), _b(name))

    # cache
    def build_cache(self):
        rc = lib().orc_build_cache(self.h, self._err, len(self._err))
        return None if rc == 0 else self._err.value.decode()

    def add_or_update_pod(self, pod: Pod):
        rc = lib().orc_add_or_update_pod(self.h, pod.h, self._err, len(self._err))
        return None if rc == 0 else self._err.value.decode()

    def remove_pod(self, pod: Pod):
        lib().orc_remove_pod(self.h, pod.h)

    def known_pod(self, uid):
        return bool(lib().orc_known_pod(self.h, _b(uid)))

    def get_node_info(self, name):
        rc = lib().orc_get_node_info(self.h, _b(name), self._err, len(self._err))
        return None if rc == 0 else self._err.value.decode()

    # filter
    def predicate(self, pod: Pod, node: str):
        self._err.value = b""
        code = lib().orc_predicate(self.h, pod.h, _b(node), self._err, len(self._err))
        return code, (None if code == PASS else self._err.value.decode())

    def filter_codes(self, pod: Pod, names: Sequence[str], names_arr=None):
        import numpy as np

        arr = names_arr if names_arr is not None else _names(names)
        out = np.empty(len(names), dtype=np.uint8)
        lib().orc_filter_codes(self.h, pod.h, arr, len(names), out.ctypes.data_as(C.POINTER(C.c_uint8)))
        return out

    def filter_json(self, pod: Pod, names: Sequence[str]) -> str:
        p = lib().orc_filter_json(self.h, pod.h, _names(names), len(names))
        s = C.string_at(p).decode()
        lib().orc_free(p)
        return s

    def filter_many(self, pods: Sequence[Pod], names: Sequence[str], threads: int = 1, names_arr=None):
        import numpy as np

        arr = names_arr if names_arr is not None else _names(names)
        ph = (C.c_void_p * len(pods))(*[p.h for p in pods])
        out = np.empty((len(pods), len(names)), dtype=np.uint8)
        lib().orc_filter_many(self.h, ph, len(pods), arr, len(names), threads,
                              out.ctypes.data_as(C.POINTER(C.c_uint8)))
        return out

    # bind
    def bind(self, pod_name, ns, uid, node, api_mode=API_OK):
        dev = C.c_int64(-1)
        self._err.value = b""
        rc = lib().orc_bind(self.h, _b(pod_name), _b(ns), _b(uid), _b(node), api_mode, C.byref(dev),
                            self._err, len(self._err))
        return dev.value, ("" if rc == 0 else self._err.value.decode())

    def allocate_gpu_id(self, node, pod: Pod):
        dev = C.c_int64(-1)
        found = lib().orc_allocate_gpu_id(self.h, _b(node), pod.h, C.byref(dev))
        return dev.value, bool(found)

    def api_pod(self, ns, name):
        h = lib().orc_api_get_pod(self.h, _b(ns), _b(name))
        return None if not h else Pod("", _h=lib().orc_pod_clone(h))

    def api_set_pod(self, pod: Pod):
        lib().orc_api_set_pod(self.h, pod.h)

    def set_now_ns(self, t):
        lib().orc_set_now_ns(self.h, int(t))

    @property
    def api_binds(self):
        return lib().orc_api_bind_count(self.h)

    # inspect
    def inspect_node(self, name, cap=64):
        total = (C.c_uint64 * cap)()
        used = (C.c_uint64 * cap)()
        n = lib().orc_inspect_node(self.h, _b(name), total, used, cap, self._err, len(self._err))
        if n < 0:
            return None
        n = min(n, cap)
        return list(total[:n]), list(used[:n])

    def bulk_load(self, node_total, node_count, node_listed, pod_node, pod_gpu, pod_mem, pod_phase) -> int:
        """Fast path for big synthetic clusters (see orc_bulk_load). Phases use the oracle's enum."""
        import numpy as np

        a = [np.ascontiguousarray(node_total, np.int64), np.ascontiguousarray(node_count, np.int64),
             np.ascontiguousarray(node_listed, np.uint8), np.ascontiguousarray(pod_node, np.int32),
             np.ascontiguousarray(pod_gpu, np.int64), np.ascontiguousarray(pod_mem, np.int64),
             np.ascontiguousarray(pod_phase, np.uint8)]
        p = [x.ctypes.data_as(C.c_void_p) for x in a]
        return lib().orc_bulk_load(self.h, len(a[0]), p[0], p[1], p[2], len(a[3]), p[3], p[4], p[5], p[6])

    # event path
    def informer_add(self, pod: Pod):
        lib().orc_informer_add(self.h, pod.h)

    def informer_update(self, old: Pod, new: Pod):
        lib().orc_informer_update(self.h, old.h, new.h)

    def informer_delete(self, pod: Pod):
        lib().orc_informer_delete(self.h, pod.h)

    def drain(self):
        rc = lib().orc_ctrl_drain(self.h, self._err, len(self._err))
        return None if rc == 0 else self._err.value.decode()

    @property
    def queue_len(self):
        return lib().orc_ctrl_queue_len(self.h)
