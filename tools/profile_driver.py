#!/usr/bin/env python
"""A short run of every kernel of the library at config C4's shapes, for ncu.

  python tools/profile_driver.py          # plain run (must exit 0 before it is profiled)
  ncu --set full --clock-control none --cache-control none --import-source on \
      -k regex:'k_expand|k_prep|k_rebuild|k_zero|k_events|k_bind_scan|k_radix|k_req_|k_node_rank|k_filter_subset' \
      -o gpurun_out/r02_kernels python tools/profile_driver.py

Twelve pipelined Filter batches into four rotating 62.7 MB outputs come first, so that by the ninth batch the L2 holds
nothing of the output being written (with --cache-control none the capture then shows the steady-state DRAM traffic);
then the compact form, a candidate-list Filter, rebuilds, pod-event batches and a bind batch.
"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

gsx = importlib.import_module("gpushare-scheduler-extender_b200")
sys.modules.setdefault("gsx", gsx)
from gsx import synth  # noqa: E402

cfg = synth.CONFIGS["C4"]
sc = synth.make_config("C4")
P, N = cfg["pending"], sc.n_nodes
reqs = synth.make_requests(P, "adversarial", cfg["seed"]).astype(np.uint64)
stride = ((N + 31) // 32 + 31) // 32 * 32
snap = gsx.Snapshot(N, sc.max_devs, sc.n_pods + 20000)
snap.node_upsert(sc.gs_nodes())
snap.pods_load(sc.gs_pod_records())
snap.rebuild()
snap.set_option(gsx.capi.OPT_PIPELINE_FILTERS, 1)
st = torch.cuda.Stream()
d_req = torch.from_numpy(reqs.view(np.int64)).cuda()
outs = [torch.empty((P, stride), dtype=torch.int32, device="cuda") for _ in range(4)]
torch.cuda.synchronize()
for i in range(12):
    snap.filter_batch_device(d_req.data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
st.synchronize()
d_rank = torch.zeros(N, dtype=torch.int16, device="cuda")
d_cut = torch.zeros(P, dtype=torch.int16, device="cuda")
for _ in range(2):
    snap.filter_batch_compact_device(d_req.data_ptr(), P, d_rank.data_ptr(), N, d_cut.data_ptr(), st.cuda_stream)
st.synchronize()
rng = np.random.default_rng(0)
snap.filter_batch_subset(reqs[:1000], rng.choice(N, 5000, replace=False).astype(np.uint32))
for _ in range(2):
    snap.rebuild()
for k in range(2):
    ev = synth.add_events(sc.n_pods + 1000 * k + np.arange(1000), rng.integers(0, N, 1000), rng.integers(0, 8, 1000),
                          rng.choice(synth.CATALOGUE[:6], 1000))
    snap.pods_apply(ev)
bits = gsx.unpack_bitmap(outs[0][:2048].cpu().numpy().view(np.uint32), N)
cnt = bits.sum(axis=1)
ok = cnt > 0
pick = np.arange(2048, dtype=np.int64) * 2654435761 % np.maximum(cnt, 1)      # a pseudo-random passing node per pod
first = (np.cumsum(bits, axis=1) > pick[:, None]).argmax(axis=1)
binds = synth.bind_records(first[ok], reqs[:2048][ok].view(np.int64), first_slot=sc.n_pods + 4000)
got = snap.bind_batch(binds)
print("profile driver ok:", int((got >= 0).sum()), "binds placed;", snap.launch_count, "launches")
snap.close()
