import importlib, os, sys, time
import numpy as np
sys.path.insert(0, "/root/repo")
gsx = importlib.import_module("gpushare-scheduler-extender_b200"); sys.modules["gsx"] = gsx
from gsx import synth
sc = synth.make_config("C4")
N=sc.n_nodes; stride=1568
snap = gsx.Snapshot(N, sc.max_devs, sc.n_pods + 1000)
snap.node_upsert(sc.gs_nodes()); snap.pods_load(sc.gs_pod_records()); snap.rebuild()
creq = gsx.PinnedBuffer(1000, np.uint64); creq.array[:] = synth.make_requests(1000, "adversarial", 0xC5).astype(np.uint64)
cout = gsx.PinnedBuffer((1000, stride), np.uint32)
snap.filter_batch(creq.array, stride, out=cout.array)
ts=[]
for i in range(200):
    time.sleep(0.01)
    a=time.perf_counter(); snap.filter_batch(creq.array, stride, out=cout.array); ts.append(time.perf_counter()-a)
ts=np.array(ts)*1e6
print("p50 %.1f p90 %.1f p99 %.1f max %.1f at %d; over 500us: %s" % (np.percentile(ts,50), np.percentile(ts,90), np.percentile(ts,99), ts.max(), ts.argmax(), np.flatnonzero(ts>500).tolist()))
