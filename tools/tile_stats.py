#!/usr/bin/env python
"""Development aid (CPU only): for a request distribution on config C4, how many of k_expand's 64-row x 128-word
tiles hold words that differ from row to row, split as the kernel splits them (<= 4 compared nodes: sparse,
more: dense).  Restates the bucket arithmetic of k_prep / k_expand phase A in numpy."""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200")
sys.modules["gsx"] = gsx
from gsx import synth  # noqa: E402

KB, TR, TW, SPARSE_MAX = 512, 64, 128, 4


def node_keys(sc):
    n, d = sc.n_nodes, sc.max_devs
    used = np.zeros((n, d), dtype=np.uint64)
    rec = sc.gs_pod_records()
    ok = (rec["dev"] >= 0) & ((rec["flags"] & 3) == 3)
    nd = sc.dev_count()[rec["node_idx"]]
    ok &= rec["dev"].astype(np.int64) < nd
    np.add.at(used, (rec["node_idx"][ok], rec["dev"][ok].astype(np.int64)), rec["mem"][ok])
    cap = sc.dev_capacity().astype(np.uint64)
    avail = cap[:, None] - used  # u64 wrap, as the reference
    mask = np.arange(d)[None, :] < sc.dev_count()[:, None]
    key = np.where(mask, avail, 0).max(axis=1)
    live = (sc.node_class() == synth.NODE_OK) & (sc.dev_count() > 0)
    return key, live, int(cap.max())


def stats(sc, req, label):
    key, live, cap_max = node_keys(sc)
    shift = 0
    while (cap_max >> shift) >= KB:
        shift += 1
    q = np.where(live, 1 + np.minimum(key >> np.uint64(shift), KB).astype(np.int64), 0)
    r = np.asarray(req, dtype=np.int64).astype(np.uint64)
    b = 1 + np.minimum(r >> np.uint64(shift), KB - 1).astype(np.int64)
    order = np.argsort(b, kind="stable")
    bs = b[order]
    n_words = (sc.n_nodes + 31) // 32
    n_cb = (n_words + TW - 1) // TW
    qpad = np.zeros(n_cb * TW * 32, dtype=np.int64)
    qpad[: sc.n_nodes] = q
    qw = qpad.reshape(n_cb * TW, 32)
    tiles = sparse_tiles = dense_tiles = sparse_words = dense_words = 0
    dense_hist = []
    for j0 in range(0, len(bs), TR):
        bmin, bmax = bs[j0], bs[min(j0 + TR, len(bs)) - 1]
        t_all = KB + 2 if bmax == KB else bmax + 1
        cmp_ = ((qw >= bmin) & (qw < t_all)).sum(axis=1).reshape(n_cb, TW)
        sp = ((cmp_ > 0) & (cmp_ <= SPARSE_MAX)).sum(axis=1)
        de = (cmp_ > SPARSE_MAX).sum(axis=1)
        tiles += n_cb
        sparse_tiles += int((sp > 0).sum())
        dense_tiles += int((de > 0).sum())
        sparse_words += int(sp.sum())
        dense_words += int(de.sum())
        dense_hist.extend(de[de > 0].tolist())
    print(f"{label:22s} shift {shift} tiles {tiles}: with sparse words {sparse_tiles} ({sparse_words} words, "
          f"{sparse_words / max(sparse_tiles, 1):.1f}/tile), with dense words {dense_tiles} ({dense_words} words, "
          f"{dense_words / max(dense_tiles, 1):.1f}/tile)")


def main():
    cfg = synth.CONFIGS["C4"]
    sc = synth.make_cluster(cfg["n_nodes"], cfg["n_devs"], cfg["dev_cap"], cfg["running"], seed=cfg["seed"])
    P = cfg["pending"]
    rng = np.random.default_rng(1)
    key, live, _ = node_keys(sc)
    print("nodes", sc.n_nodes, "live", int(live.sum()), "wrapped keys", int((key > (1 << 40)).sum()))
    hist = np.bincount((key[live & (key < (1 << 40))] >> np.uint64(9)).astype(np.int64), minlength=361)
    print("key buckets (non-empty):", {int(i): int(c) for i, c in enumerate(hist) if c})
    cases = {
        "adversarial": synth.make_requests(P, "adversarial", cfg["seed"]),
        "catalogue": synth.make_requests(P, "catalogue", cfg["seed"]),
        "mid(50k-100k)": rng.integers(50000, 100000, P),
        "top(180k-184k)": rng.integers(180000, 184320, P),
        "huge(1e6)": np.full(P, 1000000, dtype=np.int64),
    }
    for name, req in cases.items():
        stats(sc, req, name)


if __name__ == "__main__":
    main()
