#!/usr/bin/env python
"""Writes tests/golden/*.json.

The reference has no tests or fixtures and cannot be run here (no Go toolchain), so these vectors are NOT reference
outputs: they are (a) the three worked examples of the reference's design document, transcribed by hand
(docs/designs/designs.md:57,64-68,80), and (b) decisions of the CPU oracle (oracle/gpushare_oracle.c, the C
restatement of the Go code) frozen for BASELINE config 1 -- 10 nodes x 4 GPUs x 16 GiB, 100 pods through Filter + Bind.
They let the CUDA path and the oracle be checked against a fixed file, so a change to either shows up as a diff.

    python tests/golden/make_golden.py
"""
import importlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
gsx = importlib.import_module("gpushare-scheduler-extender_b200")
sys.modules["gsx"] = gsx
from gsx import synth  # noqa: E402
from oracle import oracle as O  # noqa: E402


def known_answers():
    return {
        "source": "docs/designs/designs.md (hand transcribed)",
        "ka1_filter": {"cite": "designs.md:64-68, filter.jpg", "nodes": {"N1": [16276, 12207], "N2": [12207, 12207], "N3": [8138, 16176]},
                       "dev_capacity": 16276, "request": 8138, "pass": ["N3"],
                       "failed": {"N1": "Insufficient GPU Memory in one device", "N2": "Insufficient GPU Memory in one device"}},
        "ka2_bind": {"cite": "designs.md:80, bind.jpg", "used": [4069, 8138, 12207, 0], "dev_capacity": 16276, "request": 8138, "chosen": 1},
        "ka3_capacity": {"cite": "designs.md:57", "gpu_mem": 32552, "gpu_count": 2, "per_device": 16276},
    }


def config1():
    sc = synth.make_cluster(10, 4, 16384, running=0, seed=0xC1, quirks=False)
    c = O.Cluster()
    names = [synth.node_name(i) for i in range(sc.n_nodes)]
    for i, n in enumerate(names):
        c.set_node(n, mem=int(sc.node_total[i]), count=int(sc.node_count[i]))
    rng = np.random.default_rng(0xC1)
    reqs = rng.choice(np.r_[synth.CATALOGUE[:7], [0, 16385, 6000]], 100).astype(np.int64)
    steps = []
    for i, r in enumerate(reqs):
        pod = O.Pod(f"pend-{i}", uid=f"uid-pend-{i}", limits=[int(r)])
        c.list_pod(pod)
        codes = c.filter_codes(pod, names).tolist()
        passing = [k for k, code in enumerate(codes) if code == O.PASS]
        target = passing[0] if passing else int(rng.integers(0, 10))
        dev, err = c.bind(f"pend-{i}", "default", f"uid-pend-{i}", names[target])
        steps.append({"req": int(r), "filter_codes": codes, "bind_node": target, "bind_dev": int(dev), "bind_error": err})
    final = [c.inspect_node(n)[1] for n in names]
    return {"source": "CPU oracle (C restatement of the Go handlers); not a reference output", "nodes": 10, "devices": 4,
            "dev_capacity": 16384, "steps": steps, "final_used": final}


if __name__ == "__main__":
    json.dump(known_answers(), open(os.path.join(HERE, "known_answers.json"), "w"), indent=1)
    json.dump(config1(), open(os.path.join(HERE, "config1_filter_bind.json"), "w"), indent=None, separators=(",", ":"))
    print("wrote", os.listdir(HERE))
