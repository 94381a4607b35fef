"""BASELINE config 5 in miniature (5 000 nodes instead of 50 000): batches of pod add / delete events interleaved
with Filter requests and Binds, replayed in the same logical order into the host mirror (informer handlers -> work
queue -> one device batch per drain -> CUDA kernels) and into the object-level CPU oracle (the reference's syncPod ->
AddOrUpdatePod / RemovePod, pkg/gpushare/controller.go:174-205, pkg/cache/cache.go:89-127).  After every Filter
round the two must agree on the Filter response bodies (byte for byte), on every Bind result and on the per-device
state of the nodes touched."""
import numpy as np
import pytest

from gsx import ext as X
from gsx import synth
from oracle import oracle as O

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(900)]

N_NODES, N_DEVS, CAP = 5000, 8, 184320


def test_churn_replay_linearises_like_the_reference():
    rng = np.random.default_rng(0xC5)
    names = [synth.node_name(i) for i in range(N_NODES)]
    oc = O.Cluster()
    with X.Extender(N_NODES + 8, N_DEVS, 80_000) as xe:
        for i, n in enumerate(names):
            cnt = N_DEVS if i % 97 else 4
            mem = None if i % 101 == 0 else CAP * cnt + (i % 3)
            xe.set_node(n, mem=mem, count=cnt)
            oc.set_node(n, mem=mem, count=cnt)
        serial = [0]
        live = {}                                                    # uid -> (name, kwargs) of the listed pod

        def new_pod():
            serial[0] += 1
            j = serial[0]
            node = names[int(rng.integers(0, N_NODES))]
            kw = dict(uid=f"uid-{j}", node=node, phase=int(rng.choice([X.RUNNING, X.PENDING, X.SUCCEEDED], p=[.9, .07, .03])),
                      annotations={O.IDX: int(rng.integers(0, N_DEVS + 1)) if rng.random() < 0.97 else "x",
                                   O.POD: int(rng.choice(synth.CATALOGUE[:9]))},
                      limits=[int(rng.choice(synth.CATALOGUE[:9]))])
            return f"run-{j}", kw

        # the cluster as the informers find it at start-up: BuildCache
        for _ in range(20_000):
            name, kw = new_pod()
            xe.list_pod(X.Pod(name, **kw))
            oc.list_pod(O.Pod(name, **kw))
            live[kw["uid"]] = (name, kw)
        assert xe.build_cache() is None and oc.build_cache() is None

        def event_batch(n):
            for _ in range(n):
                if rng.random() < 0.55 or not live:
                    name, kw = new_pod()
                    px, po = X.Pod(name, **kw), O.Pod(name, **kw)
                    xe.list_pod(px)
                    oc.list_pod(po)
                    xe.informer_add(px)
                    oc.informer_add(po)
                    live[kw["uid"]] = (name, kw)
                else:
                    uid = list(live)[int(rng.integers(0, len(live)))]
                    name, kw = live.pop(uid)
                    px, po = X.Pod(name, **kw), O.Pod(name, **kw)
                    xe.unlist_pod("default", name)
                    oc.unlist_pod("default", name)
                    xe.informer_delete(px)
                    oc.informer_delete(po)
            before = xe.launch_count
            assert xe.drain() == oc.drain()
            assert xe.launch_count - before <= 2                     # the whole drain is one batch on the GPU

        pend = 0
        for rnd in range(30):
            event_batch(200)
            if rnd % 5 != 4:
                continue
            touched = set()
            for k in range(6):                                       # Filter requests, one pod each
                pend += 1
                req = int(rng.choice(np.r_[synth.CATALOGUE, [0, 150000, 184320, 184321]]))
                kw = dict(uid=f"uid-pend-{pend}", limits=[req])
                px, po = X.Pod(f"pend-{pend}", **kw), O.Pod(f"pend-{pend}", **kw)
                cand = names if k % 2 == 0 else [names[i] for i in rng.choice(N_NODES, 300, replace=False)] + ["ghost"]
                assert xe.filter_json(px, cand) == oc.filter_json(po, cand), (rnd, k)
            for k in range(20):                                      # Binds, to the first node that passes
                pend += 1
                req = int(rng.choice(synth.CATALOGUE[4:]))
                kw = dict(uid=f"uid-pend-{pend}", limits=[req])
                px, po = X.Pod(f"pend-{pend}", **kw), O.Pod(f"pend-{pend}", **kw)
                xe.list_pod(px)
                oc.list_pod(po)
                sample = [names[i] for i in rng.choice(N_NODES, 64, replace=False)]
                codes = oc.filter_codes(po, sample)
                assert (xe.filter_codes([px], sample)[0] == codes).all()
                passing = np.flatnonzero(codes == O.PASS)
                target = sample[int(passing[0])] if len(passing) and rng.random() < 0.9 else sample[0]
                mode = X.API_OK if rng.random() < 0.8 else int(rng.choice([X.API_CONFLICT_THEN_OK, X.API_BIND_FAILS, X.API_OTHER_ERROR]))
                got = xe.bind(f"pend-{pend}", "default", kw["uid"], target, mode)
                assert got == oc.bind(f"pend-{pend}", "default", kw["uid"], target, mode), (rnd, k)
                touched.add(target)
            for n in list(touched) + [names[i] for i in rng.choice(N_NODES, 40, replace=False)]:
                assert xe.inspect_node(n) == oc.inspect_node(n), n
        for n in names[::7]:
            assert xe.inspect_node(n) == oc.inspect_node(n), n
