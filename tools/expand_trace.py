#!/usr/bin/env python
"""Development aid: phase timing inside k_expand.  Needs a library built with
GS_EXTRA_NVCC_FLAGS=-DGS_EXPAND_TRACE (python -m ... build --force); prints, per tile class, the mean SM cycles
between the trace stamps of thread 0.  Not part of the product or of the judged bench."""
import ctypes as C
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200")
sys.modules["gsx"] = gsx
from gsx import capi, synth  # noqa: E402

TILES, SLOTS = 6, 12
NAMES = ["wait_inputs", "phaseA+sync", "issue+ticket+request", "D_clean", "keys+compute", "D_patched", "end_sync"]


def main():
    case = os.environ.get("CASE", "adversarial")
    cfg = synth.CONFIGS["C4"]
    sc = synth.make_cluster(cfg["n_nodes"], cfg["n_devs"], cfg["dev_cap"], cfg["running"], seed=cfg["seed"])
    P, N = cfg["pending"], sc.n_nodes
    words = (N + 31) // 32
    stride = (words + 31) // 32 * 32
    snap = gsx.Snapshot(N, sc.max_devs, sc.n_pods + P)
    snap.node_upsert(sc.gs_nodes())
    snap.pods_load(sc.gs_pod_records())
    snap.rebuild()
    dev = torch.device("cuda", 0)
    st = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(st)
    outs = [torch.empty((P, stride), dtype=torch.int32, device=dev) for _ in range(4)]
    rng = np.random.default_rng(1)
    cases = {
        "adversarial": synth.make_requests(P, "adversarial", cfg["seed"]),
        "catalogue": synth.make_requests(P, "catalogue", cfg["seed"]),
        "top": rng.integers(180000, 184320, P),
        "huge": np.full(P, 1000000, dtype=np.int64),
    }
    d_req = torch.from_numpy(np.asarray(cases[case], dtype=np.int64)).to(dev)
    for i in range(6):
        snap.filter_stages_device(7, d_req.data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
    for i in range(20):  # back-to-back expand launches: the trace keeps the last one
        snap.filter_stages_device(4, d_req.data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
    torch.cuda.synchronize()
    lib = capi.load()
    n = 148 * 5 * TILES * SLOTS
    buf = np.zeros(n, dtype=np.uint64)
    rc = lib.gs_debug_expand_trace(buf.ctypes.data_as(C.POINTER(C.c_uint64)), C.c_size_t(n))
    assert rc == 0, rc
    tr = buf.reshape(148 * 5, TILES, SLOTS)
    info = tr[:, :, 8]
    # tiles actually run in the last launch: stamps of one launch are monotone per CTA
    rows = []
    t_first = int(np.median(tr[:, 0, 9]))  # globaltimer (ns) of the first tiles of the last launch
    for c in range(tr.shape[0]):
        prev_end = 0
        for t in range(TILES):
            s = tr[c, t]
            if s[0] == 0 or s[7] < s[0] or (t and s[0] < prev_end) or int(s[9]) < t_first - 5000:
                break
            prev_end = s[7]
            np_, dense, wide = int(info[c, t] & 0xffff), int((info[c, t] >> 16) & 1), int((info[c, t] >> 17) & 1)
            st_ = s[:8].astype(np.int64)
            if np_ == 0:
                st_[5] = st_[4]
                st_[6] = st_[4]
            rows.append((c, t, np_, dense, wide, int(s[9]), *np.diff(st_).tolist(), int(st_[7] - st_[0]),
                         int(s[10]) - int(s[4]), int(s[11]) - int(s[10]), int(s[5]) - int(s[11])))
    a = np.array(rows, dtype=np.int64)
    print(f"case {case}: {len(a)} tiles traced over {len(set(a[:, 0]))} CTAs")
    g0, g1 = a[:, 5].min(), a[:, 5].max()
    print(f"globaltimer span of phase-A stamps: {(g1 - g0) / 1e3:.1f} us; first tiles spread {(np.percentile(tr[:, 0, 9].astype(np.int64), 99) - g0) / 1e3:.1f} us")
    last = a[np.r_[a[1:, 0] != a[:-1, 0], True]]  # last tile of each CTA
    print("phase-A time of each CTA's last tile, us after start: p10/p50/p90/max =",
          [round(float(x - g0) / 1e3, 1) for x in np.percentile(last[:, 5], [10, 50, 90, 100])])
    per_cta = np.bincount(a[:, 0])
    print("tiles per CTA histogram:", np.bincount(per_cta).tolist())
    for label, m in (("clean", a[:, 2] == 0), ("sparse", (a[:, 2] > 0) & (a[:, 3] == 0)), ("dense", a[:, 3] == 1)):
        if not m.any():
            continue
        sel = a[m]
        print(f"{label:7s} n={len(sel):5d} mean np={sel[:, 2].mean():5.1f} | " +
              " ".join(f"{nm}={sel[:, 6 + i].mean():7.0f}" for i, nm in enumerate(NAMES)) + f" | total={sel[:, 13].mean():7.0f} cyc")
        if label == "dense":
            print(f"   dense detail: sync_after_Dclean={sel[:, 14].mean():7.0f} key_wait={sel[:, 15].mean():7.0f} search+sync={sel[:, 16].mean():7.0f}")
        for tno in range(3):
            mm = sel[sel[:, 1] == tno]
            if len(mm):
                print(f"   tile#{tno} n={len(mm):5d} " + " ".join(f"{mm[:, 6 + i].mean():7.0f}" for i in range(7)) + f" | {mm[:, 13].mean():7.0f}")


if __name__ == "__main__":
    main()
