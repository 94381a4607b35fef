"""ctypes binding of include/gpushare_b200.h.

This is plumbing for tests, bench.py and the multi-GPU driver; the product is the C-ABI library.
There is no CPU fallback: a missing library or a missing GPU raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build as _build

GS_OK, GS_ERR_INVALID, GS_ERR_CUDA, GS_ERR_CAPACITY, GS_ERR_NOMEM = 0, -1, -2, -3, -4
NODE_OK, NODE_NOT_FOUND, NODE_NOT_SHARE = 0, 1, 2
EV_ADD, EV_REMOVE = 0, 1
POD_VALID, POD_COUNTED = 1, 2
BIND_NO_DEVICE, BIND_NODE_NOT_FOUND, BIND_SLOT_BUSY = -1, -2, -3
MAX_DEVS, MAX_BATCH = 64, 12288
STAGE_SORT, STAGE_NODE_RANK, STAGE_EXPAND = 1, 2, 4
OPT_PIPELINE_FILTERS = 1
K_REBUILD, K_APPLY, K_SORT, K_NODE_RANK, K_EXPAND, K_BIND_SORT, K_BIND_SCAN, K_COMPACT, K_SUBSET = range(9)
KERNEL_NAMES = ["rebuild", "apply", "sort", "node_rank", "expand", "bind_sort", "bind_scan", "compact", "subset"]

# numpy mirrors of the C structs (layouts asserted against sizeof in tests/test_abi.py)
NODE_DT = np.dtype([("dev_capacity", "<u8"), ("node_idx", "<u4"), ("dev_count", "<u4"),
                    ("node_class", "u1"), ("pad_", "u1", (7,))])
POD_DT = np.dtype([("mem", "<u8"), ("node_idx", "<u4"), ("dev", "<i2"), ("flags", "u1"), ("pad_", "u1")])
EVENT_DT = np.dtype([("mem", "<u8"), ("slot", "<u4"), ("node_idx", "<u4"), ("dev", "<i2"), ("op", "u1"),
                     ("counted", "u1"), ("pad_", "<u4")])
BIND_DT = np.dtype([("req", "<u8"), ("commit_mem", "<u8"), ("node_idx", "<u4"), ("slot", "<u4"),
                    ("counted", "u1"), ("pad_", "u1", (7,))])
assert NODE_DT.itemsize == 24 and POD_DT.itemsize == 16 and EVENT_DT.itemsize == 24 and BIND_DT.itemsize == 32

EXPORTS = [
    "gs_last_error", "gs_abi_version", "gs_device_count", "gs_snapshot_create", "gs_snapshot_destroy",
    "gs_node_upsert", "gs_node_count", "gs_pods_load", "gs_rebuild", "gs_pods_apply", "gs_filter_batch",
    "gs_filter_batch_device", "gs_filter_stages_device", "gs_node_classes", "gs_bind_batch", "gs_bind_rollback", "gs_inspect",
    "gs_read_used", "gs_read_pods", "gs_host_alloc", "gs_host_free", "gs_profile_enable", "gs_profile_read",
    "gs_profile_reset", "gs_launch_count", "gs_device_ptrs", "gs_pods_apply_static", "gs_set_option",
    "gs_pods_apply_async", "gs_snapshot_sync", "gs_snapshot_external_write", "gs_filter_batch_subset",
    "gs_filter_batch_compact", "gs_filter_batch_compact_device", "gs_sm_count", "gs_filter_batch_compact_device_gather", "gs_debug_early_starts",
]


class GpushareError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"gpushare_b200 error {code}: {msg}")
        self.code = code


_lib = None


def library_path() -> str:
    return _build.LIB


def load():
    """dlopen the in-tree library (building it first if sources are newer). Raises if it cannot."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB
    if _build.stale():
        path = _build.build_native()
    if not os.path.exists(path):
        raise GpushareError(GS_ERR_CUDA, f"{path} is missing and could not be built; there is no CPU fallback")
    L = C.CDLL(path)
    vp, u32, u64p = C.c_void_p, C.c_uint32, C.POINTER(C.c_uint64)
    sig = {
        "gs_last_error": (C.c_char_p, []),
        "gs_abi_version": (C.c_int, []),
        "gs_device_count": (C.c_int, [C.POINTER(C.c_int)]),
        "gs_snapshot_create": (C.c_int, [C.c_int, u32, u32, u32, C.POINTER(vp)]),
        "gs_snapshot_destroy": (C.c_int, [vp]),
        "gs_node_upsert": (C.c_int, [vp, vp, u32]),
        "gs_node_count": (C.c_int, [vp, C.POINTER(u32)]),
        "gs_pods_load": (C.c_int, [vp, vp, u32, u32]),
        "gs_rebuild": (C.c_int, [vp]),
        "gs_pods_apply": (C.c_int, [vp, vp, u32]),
        "gs_filter_batch": (C.c_int, [vp, vp, u32, vp, u32]),
        "gs_filter_batch_device": (C.c_int, [vp, vp, u32, vp, u32, vp]),
        "gs_filter_stages_device": (C.c_int, [vp, u32, vp, u32, vp, u32, vp]),
        "gs_node_classes": (C.c_int, [vp, vp, u32]),
        "gs_bind_batch": (C.c_int, [vp, vp, u32, vp]),
        "gs_bind_rollback": (C.c_int, [vp, vp, u32]),
        "gs_inspect": (C.c_int, [vp, u32, vp, vp, u32, C.POINTER(u32), C.POINTER(C.c_uint8)]),
        "gs_read_used": (C.c_int, [vp, u32, u32, vp]),
        "gs_read_pods": (C.c_int, [vp, u32, u32, vp]),
        "gs_host_alloc": (C.c_int, [C.c_size_t, C.POINTER(vp)]),
        "gs_host_free": (C.c_int, [vp]),
        "gs_profile_enable": (C.c_int, [vp, C.c_int]),
        "gs_profile_read": (C.c_int, [vp, C.c_int, u64p, C.POINTER(C.c_double)]),
        "gs_profile_reset": (C.c_int, [vp]),
        "gs_launch_count": (C.c_int, [vp, u64p]),
        "gs_device_ptrs": (C.c_int, [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]),
        "gs_pods_apply_static": (C.c_int, [C.c_int, vp, u32, vp, u32, vp, u32, u32, vp, vp]),
        "gs_set_option": (C.c_int, [vp, C.c_int, C.c_int]),
        "gs_pods_apply_async": (C.c_int, [vp, vp, u32]),
        "gs_snapshot_sync": (C.c_int, [vp]),
        "gs_snapshot_external_write": (C.c_int, [vp, vp]),
        "gs_filter_batch_subset": (C.c_int, [vp, vp, u32, vp, u32, vp, u32]),
        "gs_filter_batch_compact": (C.c_int, [vp, vp, u32, vp, u32, vp]),
        "gs_filter_batch_compact_device": (C.c_int, [vp, vp, u32, vp, u32, vp, vp]),
        "gs_sm_count": (C.c_int, [vp, C.POINTER(C.c_int)]),
        "gs_debug_early_starts": (C.c_int, [vp, C.POINTER(C.c_uint64)]),
        "gs_filter_batch_compact_device_gather": (C.c_int, [vp, vp, u32, C.POINTER(vp), u32, u32, u32, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _check(rc):
    if rc != GS_OK:
        raise GpushareError(rc, load().gs_last_error().decode(errors="replace"))


def _ptr(a: np.ndarray):
    # (a.ctypes.data_as() costs ~4 us a call, more than a small Filter spends on the GPU)
    return C.c_void_p(a.__array_interface__["data"][0])


class PinnedBuffer:
    """Page-locked host memory from gs_host_alloc, viewed as a numpy array."""

    def __init__(self, shape, dtype):
        self.shape = tuple(shape) if not np.isscalar(shape) else (int(shape),)
        self.dtype = np.dtype(dtype)
        nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        p = C.c_void_p()
        _check(load().gs_host_alloc(max(nbytes, 1), C.byref(p)))
        self.ptr = p.value
        self.nbytes = nbytes
        buf = (C.c_char * max(nbytes, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def free(self):
        if self.ptr:
            self.array = None
            load().gs_host_free(C.c_void_p(self.ptr))
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Snapshot:
    """One GPU's struct-of-arrays cluster snapshot (the stand-in for cache.SchedulerCache state)."""

    def __init__(self, max_nodes: int, max_devs: int = 8, max_pods: int = 1, device: int = 0):
        L = load()
        h = C.c_void_p()
        _check(L.gs_snapshot_create(device, max_nodes, max_devs, max_pods, C.byref(h)))
        self.h = h
        self.max_nodes, self.max_devs, self.max_pods, self.device = max_nodes, max_devs, max_pods, device

    def close(self):
        if getattr(self, "h", None):
            load().gs_snapshot_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- nodes ----
    def node_upsert(self, nodes: np.ndarray):
        nodes = np.ascontiguousarray(nodes, dtype=NODE_DT)
        _check(load().gs_node_upsert(self.h, _ptr(nodes), len(nodes)))

    def upsert_nodes(self, node_idx, dev_capacity, dev_count, node_class):
        n = len(node_idx)
        a = np.zeros(n, dtype=NODE_DT)
        a["node_idx"], a["dev_capacity"], a["dev_count"], a["node_class"] = node_idx, dev_capacity, dev_count, node_class
        self.node_upsert(a)

    @property
    def node_count(self) -> int:
        v = C.c_uint32()
        _check(load().gs_node_count(self.h, C.byref(v)))
        return v.value

    @property
    def bitmap_words(self) -> int:
        return (self.node_count + 31) // 32

    def node_classes(self, n=None) -> np.ndarray:
        n = self.node_count if n is None else n
        out = np.empty(n, dtype=np.uint8)
        _check(load().gs_node_classes(self.h, _ptr(out), n))
        return out

    # ---- pods ----
    def pods_load(self, recs: np.ndarray, first_slot: int = 0):
        recs = np.ascontiguousarray(recs, dtype=POD_DT)
        _check(load().gs_pods_load(self.h, _ptr(recs), first_slot, len(recs)))

    def rebuild(self):
        _check(load().gs_rebuild(self.h))

    def pods_apply(self, events: np.ndarray):
        events = np.ascontiguousarray(events, dtype=EVENT_DT)
        _check(load().gs_pods_apply(self.h, _ptr(events), len(events)))

    def pods_apply_async(self, events: np.ndarray):
        """Queued on the GPU when this returns; every later call of the snapshot sees the batch applied."""
        events = np.ascontiguousarray(events, dtype=EVENT_DT)
        _check(load().gs_pods_apply_async(self.h, _ptr(events), len(events)))

    def sync(self):
        _check(load().gs_snapshot_sync(self.h))

    def external_write(self, stream: int = 0):
        _check(load().gs_snapshot_external_write(self.h, C.c_void_p(stream)))

    # ---- filter ----
    def filter_batch(self, req, row_stride_words: int | None = None, out: np.ndarray | None = None) -> np.ndarray:
        req = np.ascontiguousarray(req, dtype=np.uint64)
        stride = row_stride_words if row_stride_words is not None else (self.bitmap_words + 3) // 4 * 4
        if out is None:
            out = np.empty((len(req), stride), dtype=np.uint32)
        assert out.dtype == np.uint32 and out.flags.c_contiguous and out.shape == (len(req), stride)
        _check(load().gs_filter_batch(self.h, _ptr(req), len(req), _ptr(out), stride))
        return out

    def filter_batch_subset(self, req, node_subset, row_stride_words: int | None = None) -> np.ndarray:
        """[P][stride] words over the candidate list: bit c of row p = predicate(pod p, node node_subset[c])."""
        req = np.ascontiguousarray(req, dtype=np.uint64)
        sub = np.ascontiguousarray(node_subset, dtype=np.uint32)
        stride = row_stride_words if row_stride_words is not None else max(1, (len(sub) + 31) // 32)
        out = np.empty((len(req), stride), dtype=np.uint32)
        _check(load().gs_filter_batch_subset(self.h, _ptr(req), len(req), _ptr(sub), len(sub), _ptr(out), stride))
        return out

    def filter_batch_compact(self, req, rank_stride: int | None = None, out_rank: np.ndarray | None = None,
                             out_cutoff: np.ndarray | None = None):
        """(rank[chunks][stride] u16, cutoff[P] u16): pass(p, n) <=> rank[p // MAX_BATCH, n] >= cutoff[p]."""
        if not (type(req) is np.ndarray and req.dtype == np.uint64 and req.flags.c_contiguous):
            req = np.ascontiguousarray(req, dtype=np.uint64)
        n = len(req)
        if rank_stride is not None:
            stride = rank_stride
        elif out_rank is not None and out_rank.ndim == 2:
            stride = out_rank.shape[1]  # (the library checks it against the node count)
        else:
            stride = self.node_count
        chunks = max(1, (n + MAX_BATCH - 1) // MAX_BATCH)
        if out_rank is None:
            out_rank = np.empty((chunks, stride), dtype=np.uint16)
        elif not (out_rank.dtype == np.uint16 and out_rank.flags.c_contiguous and out_rank.shape == (chunks, stride)):
            raise ValueError("out_rank must be a C-contiguous uint16 array of shape (chunks, rank_stride)")
        if out_cutoff is None:
            out_cutoff = np.empty(n, dtype=np.uint16)
        elif not (out_cutoff.dtype == np.uint16 and out_cutoff.flags.c_contiguous and len(out_cutoff) == n):
            raise ValueError("out_cutoff must be a C-contiguous uint16 array with one entry per request")
        rc = _lib.gs_filter_batch_compact(self.h, _ptr(req), n, _ptr(out_rank), stride, _ptr(out_cutoff))
        if rc != GS_OK:
            _check(rc)
        return out_rank, out_cutoff

    def filter_batch_compact_device(self, d_req_ptr: int, n_pods: int, d_rank_ptr: int, rank_stride: int, d_cutoff_ptr: int,
                                    stream: int = 0):
        _check(load().gs_filter_batch_compact_device(self.h, C.c_void_p(d_req_ptr), n_pods, C.c_void_p(d_rank_ptr),
                                                     rank_stride, C.c_void_p(d_cutoff_ptr), C.c_void_p(stream)))

    def filter_batch_compact_device_gather(self, d_req_ptr: int, n_pods: int, dst_ptrs, dst_offset: int, slice_count: int,
                                           d_cutoff_ptr: int, stream: int = 0):
        """The shard's ranks written to dst_ptrs[k] + dst_offset for every k (this GPU's and the peers' gathered arrays)."""
        arr = (C.c_void_p * len(dst_ptrs))(*[C.c_void_p(int(p)) for p in dst_ptrs])
        _check(load().gs_filter_batch_compact_device_gather(self.h, C.c_void_p(d_req_ptr), n_pods, arr, len(dst_ptrs), dst_offset,
                                                            slice_count, C.c_void_p(d_cutoff_ptr), C.c_void_p(stream)))

    def filter_batch_device(self, d_req_ptr: int, n_pods: int, d_out_ptr: int, row_stride_words: int, stream: int = 0):
        _check(load().gs_filter_batch_device(self.h, C.c_void_p(d_req_ptr), n_pods, C.c_void_p(d_out_ptr),
                                             row_stride_words, C.c_void_p(stream)))

    def filter_stages_device(self, stage_mask: int, d_req_ptr: int, n_pods: int, d_out_ptr: int,
                             row_stride_words: int, stream: int = 0):
        _check(load().gs_filter_stages_device(self.h, stage_mask, C.c_void_p(d_req_ptr), n_pods, C.c_void_p(d_out_ptr),
                                              row_stride_words, C.c_void_p(stream)))

    # ---- bind ----
    def bind_batch(self, binds: np.ndarray) -> np.ndarray:
        binds = np.ascontiguousarray(binds, dtype=BIND_DT)
        out = np.empty(len(binds), dtype=np.int32)
        _check(load().gs_bind_batch(self.h, _ptr(binds), len(binds), _ptr(out)))
        return out

    def bind_rollback(self, slots):
        slots = np.ascontiguousarray(slots, dtype=np.uint32)
        _check(load().gs_bind_rollback(self.h, _ptr(slots), len(slots)))

    # ---- read-back ----
    def inspect(self, node_idx: int):
        cap = np.zeros(self.max_devs, dtype=np.uint64)
        used = np.zeros(self.max_devs, dtype=np.uint64)
        nd, cls = C.c_uint32(), C.c_uint8()
        _check(load().gs_inspect(self.h, node_idx, _ptr(cap), _ptr(used), self.max_devs, C.byref(nd), C.byref(cls)))
        return cap[: nd.value].copy(), used[: nd.value].copy(), cls.value

    def read_used(self, first_node: int = 0, n: int | None = None) -> np.ndarray:
        n = self.node_count - first_node if n is None else n
        out = np.empty((n, self.max_devs), dtype=np.uint64)
        _check(load().gs_read_used(self.h, first_node, n, _ptr(out)))
        return out

    def read_pods(self, first_slot: int, n: int) -> np.ndarray:
        out = np.empty(n, dtype=POD_DT)
        _check(load().gs_read_pods(self.h, first_slot, n, _ptr(out)))
        return out

    def set_option(self, option: int, value: int):
        _check(load().gs_set_option(self.h, option, value))

    # ---- profiling ----
    def profile(self, on: bool):
        _check(load().gs_profile_enable(self.h, 1 if on else 0))

    def profile_reset(self):
        _check(load().gs_profile_reset(self.h))

    def profile_read(self) -> dict:
        res = {}
        for kid, name in enumerate(KERNEL_NAMES):
            n, ms = C.c_uint64(), C.c_double()
            _check(load().gs_profile_read(self.h, kid, C.byref(n), C.byref(ms)))
            res[name] = (n.value, ms.value)
        return res

    def device_ptrs(self) -> dict:
        """Device addresses of the snapshot arrays (gs_device_ptrs)."""
        p = [C.c_void_p() for _ in range(4)]
        _check(load().gs_device_ptrs(self.h, *[C.byref(x) for x in p]))
        return dict(used=p[0].value, capacity=p[1].value, nodeinfo=p[2].value, pods=p[3].value)

    @property
    def sm_count(self) -> int:
        v = C.c_int()
        _check(load().gs_sm_count(self.h, C.byref(v)))
        return v.value

    @property
    def early_starts(self) -> int:
        """CTAs of the expansion kernel that found their batch's preparation unfinished (expected 0); waits for the GPU."""
        v = C.c_uint64()
        _check(load().gs_debug_early_starts(self.h, C.byref(v)))
        return v.value

    @property
    def launch_count(self) -> int:
        v = C.c_uint64()
        _check(load().gs_launch_count(self.h, C.byref(v)))
        return v.value


STATIC_MAX_SLOTS = 1 << 20


def pods_apply_static(device: int, d_events: int, n_events: int, d_pods: int, n_slots: int, d_nodeinfo: int,
                      n_nodes: int, dev_stride: int, d_used: int, stream: int = 0):
    """gs_pods_apply_static: every argument is a device address the caller owns; nothing is allocated."""
    _check(load().gs_pods_apply_static(device, C.c_void_p(d_events), n_events, C.c_void_p(d_pods), n_slots,
                                       C.c_void_p(d_nodeinfo), n_nodes, dev_stride, C.c_void_p(d_used),
                                       C.c_void_p(stream)))


def expand_compact(rank: np.ndarray, cutoff: np.ndarray, n_nodes: int, row_stride_words: int | None = None) -> np.ndarray:
    """What a caller of gs_filter_batch_compact does to get bitmap rows back (plain host arithmetic, here in
    numpy for tests and bench.py): bit n of row p = rank[p // MAX_BATCH, n] >= cutoff[p]."""
    words = (n_nodes + 31) // 32
    stride = row_stride_words if row_stride_words is not None else words
    out = np.zeros((len(cutoff), stride), dtype=np.uint32)
    for k in range(rank.shape[0]):
        lo, hi = k * MAX_BATCH, min(len(cutoff), (k + 1) * MAX_BATCH)
        bits = np.zeros((hi - lo, words * 32), dtype=np.uint8)
        bits[:, :n_nodes] = rank[k, :n_nodes][None, :] >= cutoff[lo:hi, None]
        out[lo:hi, :words] = np.packbits(bits, axis=1, bitorder="little").view(np.uint32).reshape(hi - lo, words)
    return out


def unpack_bitmap(bitmap: np.ndarray, n_nodes: int) -> np.ndarray:
    """[P][stride] uint32 words -> [P][n_nodes] bool (bit n&31 of word n>>5)."""
    b = np.unpackbits(np.ascontiguousarray(bitmap).view(np.uint8), axis=1, bitorder="little")
    return b[:, :n_nodes].astype(bool)
