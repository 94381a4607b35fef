"""ctypes front-end of oracle/fair_baseline.c ("B-fair": flat arrays, AVX2, OpenMP over all host threads).

TEST / MEASUREMENT INFRASTRUCTURE ONLY: used by bench.py's cpu_baseline legs and tests/test_fair_baseline.py.
Compiled for the CPU it runs on (-march=native), so the library is keyed by the host's CPU flags and rebuilt on
a box whose CPU differs from the one it was last built on."""
from __future__ import annotations

import ctypes as C
import hashlib
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))


def _cpu_tag() -> str:
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                return hashlib.sha1(line.encode()).hexdigest()[:10]
    except OSError:
        pass
    return "generic"


def build(force: bool = False) -> str:
    out_dir = os.path.join(_HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, f"libgpushare_fair_{_cpu_tag()}.so")
    src = os.path.join(_HERE, "fair_baseline.c")
    if force or not os.path.exists(so) or os.path.getmtime(src) > os.path.getmtime(so):
        subprocess.run(["gcc", "-O3", "-march=native", "-fopenmp", "-std=c11", "-Wall", "-Wextra", "-fPIC", "-shared",
                        "-o", so, src], check=True, capture_output=True)
    return so


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        L.fair_filter.restype = C.c_int
        L.fair_filter.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p,
                                  C.c_size_t, C.c_void_p, C.c_size_t]
        L.fair_threads.restype = C.c_int
        L.fair_has_avx2.restype = C.c_int
        _lib = L
    return _lib


def threads() -> int:
    return lib().fair_threads()


def has_avx2() -> bool:
    return bool(lib().fair_has_avx2())


def filter_bitmap(dev_capacity, used, dev_count, node_class, req, row_stride_words: int | None = None, out=None):
    """[P][stride] uint32 bitmap of Predicate.Handler over NodeInfo.Assume for every (pod, node)."""
    cap = np.ascontiguousarray(dev_capacity, dtype=np.uint64)
    used = np.ascontiguousarray(used, dtype=np.uint64)
    ndev = np.ascontiguousarray(dev_count, dtype=np.uint32)
    cls = np.ascontiguousarray(node_class, dtype=np.uint8)
    req = np.ascontiguousarray(req, dtype=np.uint64)
    n, d = used.shape
    stride = row_stride_words if row_stride_words is not None else (n + 31) // 32
    if out is None:
        out = np.empty((len(req), stride), dtype=np.uint32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rc = lib().fair_filter(p(cap), p(used), p(ndev), p(cls), n, d, p(req), len(req), p(out), stride)
    if rc != 0:
        raise MemoryError("fair_filter")
    return out
