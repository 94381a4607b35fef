"""Handler-level parity: the host mirror (C++ over the CUDA kernels) against the object-level CPU oracle, driven by
the same sequence of lister updates, cache calls, Filter requests and Binds.  Filter responses are compared as the
JSON bodies the reference would send (byte for byte), Bind results as (device, error text), state through Inspect."""
import numpy as np
import pytest

from gsx import ext as X
from oracle import oracle as O

pytestmark = pytest.mark.gpu

ODD = ["x", "-1", "", "99999999999999999999", "+2", "007"]


def both_pods(name, **kw):
    return X.Pod(name, **kw), O.Pod(name, **kw)


def check_state(xe, oc, names):
    for n in names:
        assert xe.inspect_node(n) == oc.inspect_node(n), n


def test_ka1_filter_json_through_the_mirror():
    # docs/designs/designs.md:64-68 and SURVEY.md Appendix B
    with X.Extender(8, 2, 64) as xe:
        used = {"N1": [16276, 12207], "N2": [12207, 12207], "N3": [8138, 16176]}
        for n, per in used.items():
            xe.set_node(n, mem=32552, count=2)
            for d, m in enumerate(per):
                xe.list_pod(X.Pod(f"{n}-g{d}", node=n, phase=X.RUNNING, annotations={O.IDX: d, O.POD: m}, limits=[m]))
        assert xe.build_cache() is None
        body = xe.filter_json(X.Pod("p", limits=[8138]), ["N1", "N2", "N3"])
        assert body == ('{"Nodes":null,"NodeNames":["N3"],"FailedNodes":{"N1":"Insufficient GPU Memory in one device",'
                        '"N2":"Insufficient GPU Memory in one device"},"Error":""}')


@pytest.mark.parametrize("seed", range(6))
def test_random_operation_sequences(seed):
    rng = np.random.default_rng(seed)
    n_nodes = 24
    names = [f"node-{i:03d}" for i in range(n_nodes)]
    oc = O.Cluster()
    with X.Extender(64, 8, 4096) as xe:
        def set_node(i, mem, count):
            xe.set_node(names[i], mem=mem, count=count)
            oc.set_node(names[i], mem=mem, count=count)

        for i in range(n_nodes):
            u = rng.random()
            if u < 0.1:
                set_node(i, None, 4)                      # not sharing
            elif u < 0.2:
                set_node(i, 16384 * 2, None)              # no devices
            else:
                cnt = int(rng.choice([1, 2, 4, 8]))
                set_node(i, 16384 * cnt + int(rng.integers(0, cnt)), cnt)
        live = {}
        serial = 0
        for step in range(260):
            op = rng.random()
            if op < 0.45:                                  # add / update a running pod
                serial += 1
                uid = f"U{int(rng.integers(0, 60))}"
                node = names[int(rng.integers(0, n_nodes))] if rng.random() < 0.95 else "ghost"
                idx = str(int(rng.integers(0, 9))) if rng.random() < 0.9 else str(rng.choice(ODD))
                mem = str(int(rng.choice([256, 1024, 4096, 8192, 12000]))) if rng.random() < 0.9 else str(rng.choice(ODD))
                phase = int(rng.choice([X.RUNNING, X.PENDING, X.SUCCEEDED, X.FAILED], p=[.7, .1, .1, .1]))
                kw = dict(uid=uid, node=node, phase=phase, annotations={O.IDX: idx, O.POD: mem}, limits=[1024])
                px, po = both_pods(f"p{uid}", **kw)
                assert xe.add_or_update_pod(px) == oc.add_or_update_pod(po)
                live[uid] = kw
            elif op < 0.6 and live:                        # remove (sometimes with a different device id)
                uid = str(rng.choice(sorted(live)))
                kw = dict(live[uid])
                if rng.random() < 0.2:
                    kw["annotations"] = {O.IDX: str(int(rng.integers(0, 9))), O.POD: "1"}
                px, po = both_pods(f"p{uid}", **kw)
                xe.remove_pod(px)
                oc.remove_pod(po)
                assert xe.known_pod(uid) == oc.known_pod(uid)
            elif op < 0.8:                                 # Filter
                req = int(rng.choice([0, 1, 256, 4096, 8192, 16384, 16385, -1, 2**40]))
                cand = [names[int(i)] for i in rng.integers(0, n_nodes, int(rng.integers(1, 30)))] + ["ghost"]
                px, po = both_pods(f"f{step}", limits=[req] if req else [])
                assert xe.filter_json(px, cand) == oc.filter_json(po, cand), (seed, step, req)
            elif op < 0.95:                                # Bind
                req = int(rng.choice([0, 256, 4096, 8192, 16384, 20000, -1]))
                mode = int(rng.choice([X.API_OK, X.API_OK, X.API_OK, X.API_CONFLICT_THEN_OK, X.API_OTHER_ERROR,
                                       X.API_CONFLICT_TWICE, X.API_BIND_FAILS, X.API_CONFLICT_GET_FAILS]))
                name, uid = f"b{step}", f"B{step}"
                px, po = both_pods(name, uid=uid, limits=[req] if req else [])
                xe.list_pod(px)
                oc.list_pod(po)
                xe.api_set_pod(px)
                oc.api_set_pod(po)
                node = names[int(rng.integers(0, n_nodes))] if rng.random() < 0.95 else "ghost"
                assert xe.bind(name, "default", uid, node, mode) == oc.bind(name, "default", uid, node, mode), (seed, step)
            else:                                          # capacity change / flip / node disappears and returns
                i = int(rng.integers(0, n_nodes))
                u = rng.random()
                if u < 0.3:
                    xe.delete_node(names[i])
                    oc.delete_node(names[i])
                else:
                    cnt = int(rng.choice([2, 4, 8]))
                    set_node(i, 16384 * cnt, cnt)
            if step % 20 == 0:
                check_state(xe, oc, names)
        check_state(xe, oc, names)
        pods_x = [X.Pod(f"q{i}", limits=[int(r)]) for i, r in enumerate([0, 512, 5000, 9000, 16384, 17000])]
        pods_o = [O.Pod(f"q{i}", limits=[int(r)]) for i, r in enumerate([0, 512, 5000, 9000, 16384, 17000])]
        assert (xe.filter_codes(pods_x, names + ["ghost"]) == oc.filter_many(pods_o, names + ["ghost"])).all()


def test_event_path_matches_oracle():
    """Informer events -> filter -> queue -> syncPod (pkg/gpushare/controller.go:77-100,174-332): the same random
    stream of pod creations, annotation updates (as a Bind would write them), phase changes and deletions goes to
    the mirror and to the oracle; queue lengths, knownPods and per-device used memory must agree after each drain."""
    rng = np.random.default_rng(11)
    names = [f"node-{i:02d}" for i in range(6)]
    oc = O.Cluster()
    with X.Extender(16, 4, 1024) as xe:
        for n in names:
            xe.set_node(n, mem=16384 * 4, count=4)
            oc.set_node(n, mem=16384 * 4, count=4)
        state = {}                                                # name -> kwargs of the listed pod
        for step in range(400):
            name = f"p{int(rng.integers(0, 40))}"
            old = state.get(name)
            u = rng.random()
            if old is None:
                kw = dict(uid=f"U-{name}-{step}", limits=[int(rng.choice([0, 1024, 4096]))])
                px, po = both_pods(name, **kw)
                xe.list_pod(px); oc.list_pod(po)
                xe.informer_add(px); oc.informer_add(po)
                state[name] = kw
            elif u < 0.25:                                         # deleted
                px, po = both_pods(name, **old)
                xe.unlist_pod("default", name); oc.unlist_pod("default", name)
                xe.informer_delete(px); oc.informer_delete(po)
                del state[name]
            else:                                                  # updated
                kw = dict(old)
                if u < 0.6:
                    kw.update(node=str(rng.choice(names)), annotations={O.IDX: str(int(rng.integers(-1, 5))), O.POD: str(int(rng.choice([512, 2048])))})
                elif u < 0.8:
                    kw["phase"] = int(rng.choice([X.RUNNING, X.SUCCEEDED, X.FAILED]))
                else:
                    kw["deleting"] = True
                ox, oo = both_pods(name, **old)
                nx, no = both_pods(name, **kw)
                xe.list_pod(nx); oc.list_pod(no)
                xe.informer_update(ox, nx); oc.informer_update(oo, no)
                state[name] = kw
            assert xe.queue_len == oc.queue_len
            if step % 7 == 0:
                assert xe.drain() == oc.drain()
                check_state(xe, oc, names)
                for kw in state.values():
                    assert xe.known_pod(kw["uid"]) == oc.known_pod(kw["uid"])
        assert xe.drain() == oc.drain()
        check_state(xe, oc, names)
