"""B200-native gpushare scheduling hot path: CUDA kernels behind the C ABI of include/gpushare_b200.h.

The directory name carries a hyphen, so import it with
    importlib.import_module("gpushare-scheduler-extender_b200")
(tests/conftest.py and __graft_entry__.py do exactly that and alias it as `gsx`).
"""
from . import build  # noqa: F401
from . import capi  # noqa: F401
from . import ext  # noqa: F401
from .capi import GpushareError, PinnedBuffer, Snapshot, expand_compact, unpack_bitmap  # noqa: F401

__all__ = ["build", "capi", "GpushareError", "PinnedBuffer", "Snapshot", "expand_compact", "unpack_bitmap"]
