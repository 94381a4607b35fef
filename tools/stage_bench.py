#!/usr/bin/env python
"""Development aid: device time of the K3 stages for several request distributions (CUDA events around
back-to-back launches of one stage).  Not part of the product or of the judged bench."""
import importlib
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200")
sys.modules["gsx"] = gsx
from gsx import synth  # noqa: E402


def main():
    steps = int(os.environ.get("STEPS", "50"))
    cfg = synth.CONFIGS["C4"]
    sc = synth.make_cluster(cfg["n_nodes"], cfg["n_devs"], cfg["dev_cap"], cfg["running"], seed=cfg["seed"])
    P, N = cfg["pending"], sc.n_nodes
    words = (N + 31) // 32
    stride = (words + 31) // 32 * 32
    snap = gsx.Snapshot(N, sc.max_devs, sc.n_pods + P)
    snap.node_upsert(sc.gs_nodes())
    snap.pods_load(sc.gs_pod_records())
    snap.rebuild()
    dev = torch.device("cuda", 0)
    st = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(st)
    outs = [torch.empty((P, stride), dtype=torch.int32, device=dev) for _ in range(4)]
    rng = np.random.default_rng(1)
    cases = {
        "adversarial": synth.make_requests(P, "adversarial", cfg["seed"]),
        "catalogue": synth.make_requests(P, "catalogue", cfg["seed"]),
        "small(<=1000)": rng.integers(1, 1000, P),
        "mid(50k-100k)": rng.integers(50000, 100000, P),
        "top(180k-184k)": rng.integers(180000, 184320, P),
        "too_big(>200k)": rng.integers(200000, 400000, P),
        "zero_req": np.zeros(P, dtype=np.int64),
        "big_spread(200k-260k)": rng.integers(200000, 260000, P),
        "huge(1e6)": np.full(P, 1000000, dtype=np.int64),
        "one(150000)": np.full(P, 150000, dtype=np.int64),
        "one(261000)": np.full(P, 261000, dtype=np.int64),
        "one(261631)": np.full(P, 261631, dtype=np.int64),
        "one(261632)": np.full(P, 261632, dtype=np.int64),
    }
    only = os.environ.get("CASE")
    for name, reqs in cases.items():
        if only and not name.startswith(only):
            continue
        d_req = torch.from_numpy(np.asarray(reqs, dtype=np.int64)).to(dev)
        res = {}
        for label, mask in (("pods", 1), ("nodes", 2), ("expand", 4), ("all", 7)):
            for i in range(3):
                snap.filter_stages_device(7 if i == 0 else mask, d_req.data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(steps):
                snap.filter_stages_device(mask, d_req.data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
            e1.record()
            torch.cuda.synchronize()
            res[label] = round(e0.elapsed_time(e1) / steps * 1e3, 1)
        print(f"{name:16s} us/launch: {res}", flush=True)


if __name__ == "__main__":
    main()
