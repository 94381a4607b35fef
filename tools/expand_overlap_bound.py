#!/usr/bin/env python
"""Development aid (GS_EXPAND_TRACE build): what the filter step would cost if k_prep were entirely hidden -- a run of
expand-only launches on alternating buffer sets that overlap each other (option value 2), timed with events."""
import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200")
sys.modules["gsx"] = gsx
from gsx import synth
cfg = synth.CONFIGS["C4"]
sc = synth.make_cluster(cfg["n_nodes"], cfg["n_devs"], cfg["dev_cap"], cfg["running"], seed=cfg["seed"])
P, N = cfg["pending"], sc.n_nodes
stride = ((N + 31) // 32 + 7) // 8 * 8
snap = gsx.Snapshot(N, sc.max_devs, sc.n_pods + P)
snap.node_upsert(sc.gs_nodes()); snap.pods_load(sc.gs_pod_records()); snap.rebuild()
dev = torch.device("cuda", 0)
st = torch.cuda.Stream(device=dev); torch.cuda.set_stream(st)
outs = [torch.empty((P, stride), dtype=torch.int32, device=dev) for _ in range(4)]
req = synth.make_requests(P, os.environ.get("KIND", "adversarial"), cfg["seed"], max_req=int(os.environ.get("MAXREQ", "184320")))
d_req = torch.from_numpy(np.asarray(req, dtype=np.int64)).to(dev)
for mode in ((1, 2) if os.environ.get("GS_VARIANT") == "trace" else (1,)):
    snap.set_option(1, mode)
    for i in range(4): snap.filter_batch_device(d_req.data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
    torch.cuda.synchronize()
    res = []
    for rep in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for i in range(50):
            if mode == 1: snap.filter_batch_device(d_req.data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
            else: snap.filter_stages_device(4, d_req.data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
        e1.record(st); torch.cuda.synchronize()
        res.append(e0.elapsed_time(e1) * 1e3 / 50)
    print("full pipelined steps" if mode == 1 else "expand-only overlapping", "us per launch: median %.2f min %.2f max %.2f" % (np.median(res), min(res), max(res)))
