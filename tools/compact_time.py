#!/usr/bin/env python
"""Device time of the compact Filter (three kernels) at config C4, CUDA events around loops of device-buffer calls."""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
gsx = importlib.import_module("gpushare-scheduler-extender_b200"); sys.modules.setdefault("gsx", gsx)
from gsx import synth
cfg = synth.CONFIGS["C4"]; sc = synth.make_config("C4")
P, N = cfg["pending"], sc.n_nodes
snap = gsx.Snapshot(N, sc.max_devs, sc.n_pods + 16)
snap.node_upsert(sc.gs_nodes()); snap.pods_load(sc.gs_pod_records()); snap.rebuild()
st = torch.cuda.Stream(); torch.cuda.set_stream(st)
for kind in ("adversarial", "catalogue"):
    reqs = synth.make_requests(P, kind, cfg["seed"]).astype(np.uint64)
    d_req = torch.from_numpy(reqs.view(np.int64)).cuda()
    d_rank = torch.zeros(N, dtype=torch.int16, device="cuda"); d_cut = torch.zeros(P, dtype=torch.int16, device="cuda")
    for _ in range(5):
        snap.filter_batch_compact_device(d_req.data_ptr(), P, d_rank.data_ptr(), N, d_cut.data_ptr(), st.cuda_stream)
    torch.cuda.synchronize()
    ts = []
    for rep in range(7):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            snap.filter_batch_compact_device(d_req.data_ptr(), P, d_rank.data_ptr(), N, d_cut.data_ptr(), st.cuda_stream)
        e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / 50 * 1e3)
    print(kind, "compact device time per call: median %.2f us [%.2f, %.2f]" % (sorted(ts)[3], min(ts), max(ts)))
snap.close()
