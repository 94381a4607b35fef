"""ctypes binding of include/gpushare_ext.h: the host-side mirror of the reference's cache / predicate / bind /
inspect interface (C++ inside libgpushare_b200.so).  Same shape as the Go types it stands in for."""
from __future__ import annotations

import ctypes as C
from typing import Iterable, Sequence

import numpy as np

from . import capi

PENDING, RUNNING, SUCCEEDED, FAILED, UNKNOWN = range(5)
PASS, NOT_FOUND, NOT_SHARE, INSUFFICIENT = range(4)
API_OK, API_CONFLICT_THEN_OK, API_CONFLICT_GET_FAILS, API_CONFLICT_TWICE, API_OTHER_ERROR, API_BIND_FAILS = range(6)

API_HOOK = C.CFUNCTYPE(None, C.c_void_p, C.c_char_p, C.c_char_p, C.c_int)

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    L = capi.load()
    vp, cp, i64, u64, sz, u32 = C.c_void_p, C.c_char_p, C.c_int64, C.c_uint64, C.c_size_t, C.c_uint32
    sig = {
        "gs_ext_pod_new": (vp, [cp, cp, cp]),
        "gs_ext_pod_free": (None, [vp]),
        "gs_ext_pod_set_node_name": (None, [vp, cp]),
        "gs_ext_pod_set_phase": (None, [vp, C.c_int]),
        "gs_ext_pod_set_deleting": (None, [vp, C.c_int]),
        "gs_ext_pod_set_annotation": (None, [vp, cp, cp]),
        "gs_ext_pod_add_container": (None, [vp, C.c_int, i64]),
        "gs_ext_atoi": (i64, [cp, C.POINTER(C.c_int)]),
        "gs_ext_pod_gpu_id": (i64, [vp]),
        "gs_ext_pod_annotation_mem": (u64, [vp]),
        "gs_ext_pod_resource_mem": (i64, [vp]),
        "gs_ext_create": (C.c_int, [C.c_int, u32, u32, u32, C.POINTER(vp)]),
        "gs_ext_destroy": (C.c_int, [vp]),
        "gs_ext_node_lister_set": (C.c_int, [vp, cp, C.c_int, i64, C.c_int, i64]),
        "gs_ext_node_lister_delete": (C.c_int, [vp, cp]),
        "gs_ext_pod_lister_set": (C.c_int, [vp, vp]),
        "gs_ext_pod_lister_delete": (C.c_int, [vp, cp, cp]),
        "gs_ext_api_set_pod": (C.c_int, [vp, vp]),
        "gs_ext_get_node_info": (C.c_int, [vp, cp]),
        "gs_ext_build_cache": (C.c_int, [vp]),
        "gs_ext_add_or_update_pod": (C.c_int, [vp, vp]),
        "gs_ext_remove_pod": (C.c_int, [vp, vp]),
        "gs_ext_known_pod": (C.c_int, [vp, cp]),
        "gs_ext_filter_json": (vp, [vp, vp, C.POINTER(cp), sz]),
        "gs_ext_filter_codes": (C.c_int, [vp, C.POINTER(vp), sz, C.POINTER(cp), sz, C.POINTER(C.c_uint8)]),
        "gs_ext_bind": (C.c_int, [vp, cp, cp, cp, cp, C.c_int, C.POINTER(i64), cp, sz]),
        "gs_ext_inspect_node": (C.c_int, [vp, cp, C.POINTER(u64), C.POINTER(u64), C.c_int, cp, sz]),
        "gs_ext_informer_add": (C.c_int, [vp, vp]),
        "gs_ext_informer_update": (C.c_int, [vp, vp, vp]),
        "gs_ext_informer_delete": (C.c_int, [vp, vp]),
        "gs_ext_ctrl_drain": (C.c_int, [vp]),
        "gs_ext_ctrl_queue_len": (sz, [vp]),
        "gs_ext_free": (None, [vp]),
        "gs_ext_snapshot": (vp, [vp]),
        "gs_ext_set_api_hook": (C.c_int, [vp, API_HOOK, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


EXPORTS = ["gs_ext_pod_new", "gs_ext_pod_free", "gs_ext_pod_set_node_name", "gs_ext_pod_set_phase", "gs_ext_pod_set_deleting",
           "gs_ext_pod_set_annotation", "gs_ext_pod_add_container", "gs_ext_atoi", "gs_ext_pod_gpu_id",
           "gs_ext_pod_annotation_mem", "gs_ext_pod_resource_mem", "gs_ext_create", "gs_ext_destroy",
           "gs_ext_node_lister_set", "gs_ext_node_lister_delete", "gs_ext_pod_lister_set", "gs_ext_pod_lister_delete",
           "gs_ext_api_set_pod", "gs_ext_get_node_info", "gs_ext_build_cache", "gs_ext_add_or_update_pod",
           "gs_ext_remove_pod", "gs_ext_known_pod", "gs_ext_filter_json", "gs_ext_filter_codes", "gs_ext_bind",
           "gs_ext_inspect_node", "gs_ext_free", "gs_ext_informer_add", "gs_ext_informer_update",
           "gs_ext_informer_delete", "gs_ext_ctrl_drain", "gs_ext_ctrl_queue_len", "gs_ext_snapshot",
           "gs_ext_set_api_hook"]


def _b(s: str) -> bytes:
    return s.encode("utf-8")


def _err() -> str:
    return lib().gs_last_error().decode(errors="replace")


def atoi(s: str):
    e = C.c_int(0)
    return lib().gs_ext_atoi(_b(s), C.byref(e)), e.value


class Pod:
    def __init__(self, name: str, ns: str = "default", uid: str | None = None, *, node: str = "", phase: int = PENDING,
                 deleting: bool = False, annotations: dict | None = None, limits: Iterable | None = None):
        L = lib()
        self.name, self.ns, self.uid = name, ns, uid if uid is not None else f"uid-{ns}-{name}"
        self.h = L.gs_ext_pod_new(_b(name), _b(ns), _b(self.uid))
        L.gs_ext_pod_set_node_name(self.h, _b(node))
        L.gs_ext_pod_set_phase(self.h, phase)
        L.gs_ext_pod_set_deleting(self.h, 1 if deleting else 0)
        for k, v in (annotations or {}).items():
            L.gs_ext_pod_set_annotation(self.h, _b(k), _b(str(v)))
        for lim in (limits if limits is not None else []):
            L.gs_ext_pod_add_container(self.h, 0 if lim is None else 1, 0 if lim is None else int(lim))

    def __del__(self):
        try:
            if getattr(self, "h", None):
                lib().gs_ext_pod_free(self.h)
                self.h = None
        except Exception:
            pass

    @property
    def gpu_id(self):
        return lib().gs_ext_pod_gpu_id(self.h)

    @property
    def annotation_mem(self):
        return lib().gs_ext_pod_annotation_mem(self.h)

    @property
    def resource_mem(self):
        return lib().gs_ext_pod_resource_mem(self.h)


class Extender:
    """Listers + cache mirror + handlers over one GPU snapshot."""

    def __init__(self, max_nodes: int, max_devs: int = 8, max_pods: int = 1024, device: int = 0):
        h = C.c_void_p()
        if lib().gs_ext_create(device, max_nodes, max_devs, max_pods, C.byref(h)) != 0:
            raise capi.GpushareError(-2, _err())
        self.h = h
        self._buf = C.create_string_buffer(2048)

    def close(self):
        if getattr(self, "h", None):
            lib().gs_ext_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_node(self, name, mem=None, count=None):
        lib().gs_ext_node_lister_set(self.h, _b(name), mem is not None, int(mem or 0), count is not None, int(count or 0))

    def delete_node(self, name):
        lib().gs_ext_node_lister_delete(self.h, _b(name))

    def list_pod(self, pod: Pod):
        lib().gs_ext_pod_lister_set(self.h, pod.h)

    def unlist_pod(self, ns, name):
        lib().gs_ext_pod_lister_delete(self.h, _b(ns), _b(name))

    def api_set_pod(self, pod: Pod):
        lib().gs_ext_api_set_pod(self.h, pod.h)

    def get_node_info(self, name):
        return None if lib().gs_ext_get_node_info(self.h, _b(name)) == 0 else _err()

    def build_cache(self):
        return None if lib().gs_ext_build_cache(self.h) == 0 else _err()

    def add_or_update_pod(self, pod: Pod):
        return None if lib().gs_ext_add_or_update_pod(self.h, pod.h) == 0 else _err()

    def remove_pod(self, pod: Pod):
        lib().gs_ext_remove_pod(self.h, pod.h)

    def known_pod(self, uid):
        return bool(lib().gs_ext_known_pod(self.h, _b(uid)))

    def filter_json(self, pod: Pod, names: Sequence[str]) -> str:
        arr = (C.c_char_p * len(names))(*[_b(n) for n in names])
        p = lib().gs_ext_filter_json(self.h, pod.h, arr, len(names))
        if not p:
            raise capi.GpushareError(-2, _err())
        s = C.string_at(p).decode()
        lib().gs_ext_free(p)
        return s

    def filter_codes(self, pods: Sequence[Pod], names: Sequence[str]) -> np.ndarray:
        arr = (C.c_char_p * len(names))(*[_b(n) for n in names])
        ph = (C.c_void_p * len(pods))(*[p.h for p in pods])
        out = np.empty((len(pods), len(names)), dtype=np.uint8)
        if lib().gs_ext_filter_codes(self.h, ph, len(pods), arr, len(names), out.ctypes.data_as(C.POINTER(C.c_uint8))) != 0:
            raise capi.GpushareError(-2, _err())
        return out

    def bind(self, pod_name, ns, uid, node, api_mode=API_OK):
        dev = C.c_int64(-1)
        buf = C.create_string_buffer(2048)  # per call: binds may come from several threads
        rc = lib().gs_ext_bind(self.h, _b(pod_name), _b(ns), _b(uid), _b(node), api_mode, C.byref(dev), buf, len(buf))
        return dev.value, ("" if rc == 0 else buf.value.decode())

    def set_api_hook(self, fn):
        """fn(pod_key: str, node: str, api_mode: int) runs where Allocate's apiserver round trips happen (None = off)."""
        if fn is None:
            self._hook = API_HOOK(0)
        else:
            self._hook = API_HOOK(lambda user, key, node, mode: fn(key.decode(), node.decode(), mode))
        lib().gs_ext_set_api_hook(self.h, self._hook, None)

    @property
    def launch_count(self) -> int:
        """Kernels launched so far by the snapshot underneath."""
        v = C.c_uint64()
        capi._check(capi.load().gs_launch_count(C.c_void_p(lib().gs_ext_snapshot(self.h)), C.byref(v)))
        return v.value

    def informer_add(self, pod: Pod):
        lib().gs_ext_informer_add(self.h, pod.h)

    def informer_update(self, old: Pod, new: Pod):
        lib().gs_ext_informer_update(self.h, old.h, new.h)

    def informer_delete(self, pod: Pod):
        lib().gs_ext_informer_delete(self.h, pod.h)

    def drain(self):
        return None if lib().gs_ext_ctrl_drain(self.h) == 0 else _err()

    @property
    def queue_len(self):
        return lib().gs_ext_ctrl_queue_len(self.h)

    def inspect_node(self, name, cap=64):
        total, used = (C.c_uint64 * cap)(), (C.c_uint64 * cap)()
        n = lib().gs_ext_inspect_node(self.h, _b(name), total, used, cap, self._buf, len(self._buf))
        if n < 0:
            return None
        n = min(n, cap)
        return list(total[:n]), list(used[:n])
