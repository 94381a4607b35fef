#!/usr/bin/env python
"""Where the time of one gs_filter_batch_compact call with page-locked buffers goes: the enqueue of its three kernels
(device-buffer entry point on the same buffers, no wait), the kernels themselves (events), the whole synchronous call."""
import importlib, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200")
sys.modules["gsx"] = gsx
from gsx import synth
cfg = synth.CONFIGS["C4"]
sc = synth.make_config("C4")
P, N = cfg["pending"], sc.n_nodes
snap = gsx.Snapshot(N, sc.max_devs, sc.n_pods + 1000)
snap.node_upsert(sc.gs_nodes()); snap.pods_load(sc.gs_pod_records()); snap.rebuild()
reqs = synth.make_requests(P, "adversarial", cfg["seed"]).astype(np.uint64)
h_req = gsx.PinnedBuffer(P, np.uint64); h_req.array[:] = reqs
h_rank = gsx.PinnedBuffer((1, N), np.uint16); h_cut = gsx.PinnedBuffer(P, np.uint16)
st = torch.cuda.Stream()
def med(f, n=300, w=30):
    for _ in range(w): f()
    ts = []
    for _ in range(n):
        a = time.perf_counter(); f(); ts.append(time.perf_counter() - a)
    return 1e6 * float(np.median(ts)), 1e6 * float(np.percentile(ts, 99))
print("whole call (page-locked buffers): p50 %.1f us p99 %.1f us" % med(lambda: snap.filter_batch_compact(h_req.array, out_rank=h_rank.array, out_cutoff=h_cut.array)))
def enqueue():
    snap.filter_batch_compact_device(h_req.ptr, P, h_rank.ptr, N, h_cut.ptr, st.cuda_stream)
def enq_sync():
    enqueue(); st.synchronize()
print("enqueue + stream synchronize:     p50 %.1f us p99 %.1f us" % med(enq_sync))
ts = []
for _ in range(200):
    st.synchronize(); a = time.perf_counter(); enqueue(); ts.append(time.perf_counter() - a)
st.synchronize()
print("enqueue only (three launches):    p50 %.1f us" % (1e6 * float(np.median(ts))))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
dev = []
for _ in range(100):
    e0.record(st); enqueue(); e1.record(st); st.synchronize(); dev.append(e0.elapsed_time(e1) * 1e3)
print("device time of the three kernels: p50 %.1f us" % float(np.median(dev)))
d_req = torch.from_numpy(reqs.view(np.int64)).cuda(); d_rank = torch.zeros(N, dtype=torch.int16, device="cuda"); d_cut = torch.zeros(P, dtype=torch.int16, device="cuda")
dev = []
for _ in range(100):
    e0.record(st); snap.filter_batch_compact_device(d_req.data_ptr(), P, d_rank.data_ptr(), N, d_cut.data_ptr(), st.cuda_stream); e1.record(st); st.synchronize(); dev.append(e0.elapsed_time(e1) * 1e3)
print("the same with device buffers:     p50 %.1f us" % float(np.median(dev)))
for name, (rq, rk, ct) in {"requests in host memory, results in device memory": (h_req.ptr, d_rank.data_ptr(), d_cut.data_ptr()),
                           "requests in device memory, results in host memory": (d_req.data_ptr(), h_rank.ptr, h_cut.ptr),
                           "only rank[] in host memory": (d_req.data_ptr(), h_rank.ptr, d_cut.data_ptr()),
                           "only cutoff[] in host memory": (d_req.data_ptr(), d_rank.data_ptr(), h_cut.ptr)}.items():
    dev = []
    for _ in range(100):
        e0.record(st); snap.filter_batch_compact_device(rq, P, rk, N, ct, st.cuda_stream); e1.record(st); st.synchronize(); dev.append(e0.elapsed_time(e1) * 1e3)
    print("%-52s p50 %.1f us" % (name + ":", float(np.median(dev))))
