#!/usr/bin/env python
"""Development aid (GS_EXPAND_TRACE build): timeline of the k_prep / k_expand launches of a run of pipelined filter
calls: per launch, when its first / last block entered, passed its start wait, and left (globaltimer, us)."""
import ctypes as C, importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200")
sys.modules["gsx"] = gsx
from gsx import capi, synth
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
req = synth.make_requests(P, os.environ.get("KIND", "catalogue"), cfg["seed"])
d_req = torch.from_numpy(np.asarray(req, dtype=np.int64)).to(dev)
lib = capi.load()
for i in range(10): snap.filter_batch_device(d_req.data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
torch.cuda.synchronize()
assert lib.gs_debug_ktrace_reset() == 0
nsteps = 24
for i in range(nsteps): snap.filter_batch_device(d_req.data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
torch.cuda.synchronize()
lo = np.zeros((64, 2, 6), dtype=np.uint64); hi = np.zeros((64, 2, 6), dtype=np.uint64)
assert lib.gs_debug_ktrace_read(lo.ctypes.data_as(C.POINTER(C.c_uint64)), hi.ctypes.data_as(C.POINTER(C.c_uint64))) == 0
valid = [q for q in range(64) if hi[q, 1, 0] != 0]
order = sorted(valid, key=lambda q: int(lo[q, 1, 0]))
t0 = int(lo[order[0], 0, 0])
f = lambda v: f"{(int(v) - t0) / 1e3:7.1f}"
print("launch | prep: entry first..last, [atomic back first..last, counts complete first..last], past-wait last, exit last | expand: entry first..last, past-wait first..last, exit first..last")
for q in order[4:16]:
    print(f"{q:4d}   | {f(lo[q,0,0])}..{f(hi[q,0,0])} [{f(lo[q,0,3])}..{f(hi[q,0,3])} {f(lo[q,0,4])}..{f(hi[q,0,4])}] {f(hi[q,0,1])} {f(hi[q,0,2])} | {f(lo[q,1,0])}..{f(hi[q,1,0])}  {f(lo[q,1,1])}..{f(hi[q,1,1])}  {f(lo[q,1,2])}..{f(hi[q,1,2])}")
