"""Committed golden vectors (tests/golden/, made by tests/golden/make_golden.py): the design-document examples and
the frozen config-1 decisions.  CPU: the oracle still reproduces them.  GPU: the CUDA path reproduces them."""
import json
import os

import numpy as np
import pytest

import gsx
from gsx import synth
from oracle import oracle as O

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
KA = json.load(open(os.path.join(G, "known_answers.json")))
C1 = json.load(open(os.path.join(G, "config1_filter_bind.json")))
NAMES = [synth.node_name(i) for i in range(C1["nodes"])]


def test_oracle_reproduces_config1_golden():
    c = O.Cluster()
    for n in NAMES:
        c.set_node(n, mem=C1["dev_capacity"] * C1["devices"], count=C1["devices"])
    for i, st in enumerate(C1["steps"]):
        pod = O.Pod(f"pend-{i}", uid=f"uid-pend-{i}", limits=[st["req"]])
        c.list_pod(pod)
        assert c.filter_codes(pod, NAMES).tolist() == st["filter_codes"]
        assert c.bind(f"pend-{i}", "default", f"uid-pend-{i}", NAMES[st["bind_node"]]) == (st["bind_dev"], st["bind_error"])
    assert [c.inspect_node(n)[1] for n in NAMES] == C1["final_used"]


def test_oracle_reproduces_design_doc_examples():
    k = KA["ka1_filter"]
    c = O.Cluster()
    for n, per in k["nodes"].items():
        c.set_node(n, mem=k["dev_capacity"] * 2, count=2)
        for d, m in enumerate(per):
            c.add_or_update_pod(O.Pod(f"{n}{d}", node=n, phase=O.RUNNING, annotations={O.IDX: d, O.POD: m}))
    body = json.loads(c.filter_json(O.Pod("p", limits=[k["request"]]), list(k["nodes"])))
    assert body["NodeNames"] == k["pass"] and body["FailedNodes"] == k["failed"]
    assert KA["ka3_capacity"]["gpu_mem"] // KA["ka3_capacity"]["gpu_count"] == KA["ka3_capacity"]["per_device"]


@pytest.mark.gpu
def test_cuda_reproduces_config1_golden():
    with gsx.Snapshot(C1["nodes"], C1["devices"], 256) as s:
        s.upsert_nodes(range(C1["nodes"]), [C1["dev_capacity"]] * C1["nodes"], [C1["devices"]] * C1["nodes"], [0] * C1["nodes"])
        for i, st in enumerate(C1["steps"]):
            bits = gsx.unpack_bitmap(s.filter_batch([np.uint64(st["req"])]), C1["nodes"])[0]
            assert np.where(bits, O.PASS, O.INSUFFICIENT).tolist() == st["filter_codes"]
            dev = int(s.bind_batch(synth.bind_records([st["bind_node"]], [st["req"]], first_slot=i))[0])
            assert dev == st["bind_dev"] and (dev >= 0) == (st["bind_error"] == "")
        assert s.read_used().tolist() == C1["final_used"]


@pytest.mark.gpu
def test_cuda_reproduces_design_doc_bind_example():
    k = KA["ka2_bind"]
    with gsx.Snapshot(1, 4, 8) as s:
        s.upsert_nodes([0], [k["dev_capacity"]], [4], [0])
        live = [(d, m) for d, m in enumerate(k["used"]) if m]
        s.pods_apply(synth.add_events(range(len(live)), [0] * len(live), [d for d, _ in live], [m for _, m in live]))
        assert int(s.bind_batch(synth.bind_records([0], [k["request"]], first_slot=7))[0]) == k["chosen"]
