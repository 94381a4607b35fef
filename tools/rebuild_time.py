#!/usr/bin/env python
"""Device time of gs_rebuild (K1: zeroing + k_rebuild) at config C4, for a pod table in random slot order (the synthetic
cluster's) and for the same table with slots grouped by node (what a host that allocates a node's slots together gets)."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200"); sys.modules["gsx"] = gsx
from gsx import synth
sc = synth.make_config("C4")
recs = sc.gs_pod_records()
orders = (("random slot order", recs), ("slots grouped by node", recs[np.argsort(recs["node_idx"], kind="stable")]),
          ("slots sorted by (node, device)", recs[np.lexsort((recs["dev"], recs["node_idx"]))]))
for label, r in orders:
    snap = gsx.Snapshot(sc.n_nodes, sc.max_devs, sc.n_pods + 10)
    snap.node_upsert(sc.gs_nodes()); snap.pods_load(r); snap.rebuild()
    u0 = snap.read_used().copy()
    snap.profile(True); snap.profile_reset()
    for _ in range(20): snap.rebuild()
    pr = snap.profile_read()["rebuild"]
    us = pr[1] / (pr[0] // 2) * 1e3
    nbytes = (sc.n_pods + 10) * 16 + sc.n_nodes * sc.max_devs * 8
    print(f"{label}: {us:.2f} us per call (zeroing + k_rebuild, CUDA events) = {nbytes / us / 1e3:.0f} GB/s; same result: {bool((snap.read_used() == u0).all())}")
    snap.close()
