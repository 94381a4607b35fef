#!/usr/bin/env python
"""Where a gs_bind_batch call spends its time (C4 snapshot): host wall time of the call and the device time of its two
kernels (the library's per-kernel counters), for a few batch sizes."""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200"); sys.modules["gsx"] = gsx
from gsx import synth
sc = synth.make_config("C4")
snap = gsx.Snapshot(sc.n_nodes, sc.max_devs, sc.n_pods + 200000)
snap.node_upsert(sc.gs_nodes()); snap.pods_load(sc.gs_pod_records()); snap.rebuild()
rng = np.random.default_rng(0)
slot = sc.n_pods
for B in (1, 16, 200, 2048):
    for prof in (False, True):
        snap.profile(prof); snap.profile_reset()
        ts = []
        for it in range(30):
            binds = synth.bind_records(rng.integers(0, sc.n_nodes, B), rng.choice(synth.CATALOGUE[:5], B), first_slot=slot)
            slot += B
            t0 = time.perf_counter(); snap.bind_batch(binds); ts.append(time.perf_counter() - t0)
        ts = sorted(ts[5:])
        line = f"B={B} profiling={prof}: wall p50 {1e6*ts[len(ts)//2]:.1f} us"
        if prof:
            pr = snap.profile_read()
            per = lambda k: f"{1e3*pr[k][1]/pr[k][0]:.1f} us" if pr[k][0] else "-"   # (a queue of <= 32 binds is one launch)
            line += f"; device: sort {per('bind_sort')}, scan {per('bind_scan')}"
        print(line)
snap.close()
