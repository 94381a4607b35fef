#!/usr/bin/env python
"""Development aid: host time to issue pipelined filter calls vs device time per call."""
import importlib, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200")
sys.modules["gsx"] = gsx
from gsx import synth
cfg = synth.CONFIGS["C4"]
sc = synth.make_cluster(cfg["n_nodes"], cfg["n_devs"], cfg["dev_cap"], cfg["running"], seed=cfg["seed"])
P, N = cfg["pending"], sc.n_nodes
words = (N + 31) // 32
stride = (words + 7) // 8 * 8
snap = gsx.Snapshot(N, sc.max_devs, sc.n_pods + P)
snap.node_upsert(sc.gs_nodes()); snap.pods_load(sc.gs_pod_records()); snap.rebuild()
snap.set_option(1, int(os.environ.get("PIPE", "1")))
dev = torch.device("cuda", 0)
st = torch.cuda.Stream(device=dev); torch.cuda.set_stream(st)
outs = [torch.empty((P, stride), dtype=torch.int32, device=dev) for _ in range(4)]
req = synth.make_requests(P, os.environ.get("KIND", "adversarial"), cfg["seed"])
d_req = torch.from_numpy(np.asarray(req, dtype=np.int64)).to(dev)
ptrs = [o.data_ptr() for o in outs]; rp = d_req.data_ptr(); sh = st.cuda_stream
for i in range(20): snap.filter_batch_device(rp, P, ptrs[i % 4], stride, sh)
torch.cuda.synchronize()
if os.environ.get("EXPAND_ONLY"):
    snap.set_option(1, 2)
    call = lambda i: snap.filter_stages_device(4, rp, P, ptrs[i % 4], stride, sh)
else:
    call = lambda i: snap.filter_batch_device(rp, P, ptrs[i % 4], stride, sh)
for steps in [int(x) for x in os.environ.get("STEPS", "200,1000").split(",")]:
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for i in range(steps): call(i)
    t1 = time.perf_counter(); e1.record(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"steps {steps}: host issue {1e6*(t1-t0)/steps:.1f} us/call, device {1e3*e0.elapsed_time(e1)/steps:.1f} us/call, wall {1e6*(t2-t0)/steps:.1f}")
