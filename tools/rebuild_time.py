import importlib, sys, os
import numpy as np, torch
sys.path.insert(0, "/root/repo")
gsx = importlib.import_module("gpushare-scheduler-extender_b200"); sys.modules["gsx"] = gsx
from gsx import synth
cfg = synth.CONFIGS["C4"]
sc = synth.make_cluster(cfg["n_nodes"], cfg["n_devs"], cfg["dev_cap"], cfg["running"], seed=cfg["seed"])
snap = gsx.Snapshot(sc.n_nodes, sc.max_devs, sc.n_pods + 10)
snap.node_upsert(sc.gs_nodes()); snap.pods_load(sc.gs_pod_records()); snap.rebuild()
u0 = snap.read_used().copy()
snap.profile(True); snap.profile_reset()
for _ in range(20): snap.rebuild()
pr = snap.profile_read()["rebuild"]
print("rebuild per call (memset + kernel, device events):", round(pr[1]/pr[0]*1e3, 2), "us", "same result:", bool((snap.read_used()==u0).all()))
