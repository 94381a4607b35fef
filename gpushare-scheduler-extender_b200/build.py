"""Builds libgpushare_b200.so in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo)."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
# development builds: GS_VARIANT=<name> builds and loads lib/libgpushare_b200_<name>.so (with GS_EXTRA_NVCC_FLAGS), so that
# several compile-time variants of a kernel can be timed in one GPU session; the product is the unsuffixed library
_VARIANT = os.environ.get("GS_VARIANT", "")
LIB_DIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIB_DIR, "libgpushare_b200%s.so" % ("_" + _VARIANT if _VARIANT else ""))

SOURCES = ["api.cu", "filter_kernels.cu", "state_kernels.cu", "compact_kernels.cu", "noalloc.cu", "ext.cpp"]
HEADERS = ["common.cuh", "kernels.h", "state_device.cuh", "exports.map", "ext_state.inc", "ext_handlers.inc",
           os.path.join(ROOT, "include", "gpushare_ext.h"), os.path.join(ROOT, "include", "gpushare_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC,-Wall",
    "-I", os.path.join(ROOT, "include"), "-I", CSRC,
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: the gpushare B200 path has no CPU fallback and cannot be built without it")


def stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + [h if os.path.isabs(h) else os.path.join(CSRC, h) for h in HEADERS]
    return any(os.path.getmtime(d) > t for d in deps)


def build_native(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        return LIB
    os.makedirs(LIB_DIR, exist_ok=True)
    nvcc = _nvcc()
    extra = os.environ.get("GS_EXTRA_NVCC_FLAGS", "").split()  # development builds only (e.g. -DGS_EXPAND_TRACE)
    objs = []
    for src in SOURCES:
        obj = os.path.join(LIB_DIR, os.path.splitext(src)[0] + ("_" + _VARIANT if _VARIANT else "") + ".o")
        cmd = [nvcc, *NVCC_FLAGS, *extra, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if verbose:
            print(r.stderr)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        objs.append(obj)
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-Xlinker",
           "--version-script=" + os.path.join(CSRC, "exports.map"), "-o", LIB, *objs]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    import sys

    print(build_native(force="--force" in sys.argv, verbose="-v" in sys.argv))
