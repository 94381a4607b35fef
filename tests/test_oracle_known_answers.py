"""Pins the CPU oracle against the only known-answer material the reference ships: the worked
examples of docs/designs/designs.md (KA-1 :64-68 + filter.jpg, KA-2 :80 + bind.jpg, KA-3 :57).
The reference has no tests, so beyond these three vectors parity is unpinned."""
import json

from oracle import oracle as O


def _running(name, node, idx, mem):
    return O.Pod(name, node=node, phase=O.RUNNING, annotations={O.IDX: idx, O.POD: mem}, limits=[mem])


def test_ka3_capacity_report():
    # designs.md:57 -- 2 GPUs x 16276 MiB -> gpu-mem 32552, gpu-count 2 -> 16276 per device
    c = O.Cluster()
    c.set_node("N1", mem=16276 * 2, count=2)
    total, used = c.inspect_node("N1")
    assert total == [16276, 16276] and used == [0, 0]


def test_ka1_filter_example():
    # designs.md:64-68: 3 nodes x 2 GPUs x 16276; request 8138 -> only N3 passes
    c = O.Cluster()
    used = {"N1": [16276, 12207], "N2": [12207, 12207], "N3": [8138, 16176]}
    for n, per_dev in used.items():
        c.set_node(n, mem=32552, count=2)
        for d, m in enumerate(per_dev):
            p = _running(f"{n}-g{d}", n, d, m)
            c.list_pod(p)
    assert c.build_cache() is None
    for n, per_dev in used.items():
        total, u = c.inspect_node(n)
        assert u == per_dev and total == [16276, 16276]
    pod = O.Pod("p", limits=[8138])
    codes = c.filter_codes(pod, ["N1", "N2", "N3"])
    assert list(codes) == [O.INSUFFICIENT, O.INSUFFICIENT, O.PASS]
    body = c.filter_json(pod, ["N1", "N2", "N3"])
    # byte-for-byte shape of the reference response for KA-1
    assert body == ('{"Nodes":null,"NodeNames":["N3"],"FailedNodes":{"N1":"Insufficient GPU Memory in one device",'
                    '"N2":"Insufficient GPU Memory in one device"},"Error":""}')
    assert json.loads(body)["NodeNames"] == ["N3"]


def test_ka2_bind_example():
    # designs.md:80: one node x 4 GPUs x 16276; used {4069, 8138, 12207, 0}; request 8138 -> GPU1
    c = O.Cluster()
    c.set_node("N1", mem=16276 * 4, count=4)
    for d, m in enumerate([4069, 8138, 12207, 0]):
        if m:
            c.list_pod(_running(f"r{d}", "N1", d, m))
    assert c.build_cache() is None
    pod = O.Pod("p", uid="u-p", limits=[8138])
    c.list_pod(pod)
    dev, found = c.allocate_gpu_id("N1", pod)
    assert (dev, found) == (1, True)
    c.set_now_ns(1234567890)
    dev, err = c.bind("p", "default", "u-p", "N1")
    assert err == "" and dev == 1
    total, used = c.inspect_node("N1")
    assert used == [4069, 8138 + 8138, 12207, 0]
    ap = c.api_pod("default", "p")
    assert ap.annotation(O.IDX) == "1" and ap.annotation(O.POD) == "8138"
    assert ap.annotation(O.DEV) == "16276" and ap.annotation(O.ASSIGNED) == "false"
    assert ap.annotation(O.ASSUME_TIME) == "1234567890"
    assert c.api_binds == 1
