"""The numpy struct-of-arrays reference (oracle/soa_reference.py) against the faithful C oracle.
CPU only.  This is what licenses using the numpy reference as the checker at sizes the object-level
oracle cannot reach."""
import numpy as np
import pytest

from gsx import synth
from oracle import oracle as O
from oracle import soa_reference as S
from tests import helpers as H


@pytest.mark.parametrize("seed", range(12))
def test_filter_and_inspect_match_oracle(seed):
    sc = synth.make_cluster(n_nodes=40, n_devs=4, dev_cap=16384, running=260, seed=seed, quirks=True, fill=0.7)
    c = H.oracle_cluster_tolerant(sc)
    cap, used, ndev, cls, _ = H.soa_state(sc)
    names = H.all_names(sc)
    for i in range(sc.n_nodes):
        got = c.inspect_node(names[i])
        if not sc.node_listed[i]:
            assert got is None
            continue
        total, u = got
        assert total == [int(cap[i])] * int(ndev[i])
        assert u == [int(x) for x in used[i, : int(ndev[i])]]
    reqs = np.r_[synth.make_requests(24, "adversarial", seed, max_req=20000), [0, -1, 16384, 16385, 2**40]].astype(np.int64)
    key, elig = S.node_keys(cap, used, ndev, cls)
    want = S.filter_codes(reqs.astype(np.uint64), key, elig, cls)
    for p, r in enumerate(reqs):
        got = c.filter_codes(H.pending_pod(p, r), names)
        assert (got == want[p]).all(), (seed, p, int(r))


@pytest.mark.parametrize("seed", range(6))
def test_bind_sequence_matches_oracle(seed):
    sc = synth.make_cluster(n_nodes=12, n_devs=4, dev_cap=16384, running=60, seed=100 + seed, quirks=True, fill=0.6)
    c = H.oracle_cluster_tolerant(sc)
    cap, used, ndev, cls, recs = H.soa_state(sc)
    rng = np.random.default_rng(seed)
    B = 120
    nodes = rng.integers(0, sc.n_nodes, B)
    reqs = rng.choice(np.r_[synth.CATALOGUE[:7], [0, 3000, 5000, 16384]], B).astype(np.int64)
    binds = synth.bind_records(nodes, reqs, first_slot=sc.n_pods)
    table = np.zeros(sc.n_pods + B, dtype=recs.dtype)
    table[: sc.n_pods] = recs
    want = S.bind_sequential(binds, cap, used, ndev, cls, table)
    names = H.all_names(sc)
    for i in range(B):
        pod = H.pending_pod(i, reqs[i])
        c.list_pod(pod)
        dev, err = c.bind(f"pend-{i}", "default", f"uid-pend-{i}", names[nodes[i]])
        if want[i] == -2:
            assert err == f'node "{names[nodes[i]]}" not found'
        elif want[i] == -1:
            assert dev == -1 and err == f"The node  can't place the pod pend-{i} in ns default"
        else:
            assert (dev, err) == (int(want[i]), "")
    for i in range(sc.n_nodes):
        if sc.node_listed[i]:
            assert c.inspect_node(names[i])[1] == [int(x) for x in used[i, : int(ndev[i])]]


def test_events_match_oracle():
    sc = synth.make_cluster(n_nodes=8, n_devs=4, dev_cap=16384, running=0, seed=5, quirks=False)
    c = H.oracle_cluster_tolerant(sc)
    cap, used, ndev, cls, _ = H.soa_state(sc)
    names = H.all_names(sc)
    rng = np.random.default_rng(7)
    n_uid = 40
    table = np.zeros(n_uid, dtype=synth.POD_DT)
    where = {}
    evs = []
    for step in range(400):
        u = int(rng.integers(0, n_uid))
        if u in where and rng.random() < 0.45:
            node, dev, mem, ph = where.pop(u)
            p = O.Pod(f"p{u}", uid=f"U{u}", node=names[node], phase=ph, annotations={O.IDX: dev, O.POD: mem})
            c.remove_pod(p)
            evs.append(synth.remove_events([u]))
        else:
            # a UID keeps its (node, device) for life here: one slot per UID is then exact
            node, dev = where[u][:2] if u in where else (int(rng.integers(0, 8)), int(rng.integers(0, 4)))
            mem = int(rng.choice(synth.CATALOGUE[:6]))
            ph = int(rng.choice([O.RUNNING, O.PENDING, O.SUCCEEDED, O.FAILED], p=[.7, .1, .1, .1]))
            where[u] = (node, dev, mem, ph)
            p = O.Pod(f"p{u}", uid=f"U{u}", node=names[node], phase=ph, annotations={O.IDX: dev, O.POD: mem})
            assert c.add_or_update_pod(p) is None
            evs.append(synth.add_events([u], [node], [dev], [mem], counted=ph not in (O.SUCCEEDED, O.FAILED)))
    S.apply_events(np.concatenate(evs), table, used, ndev)
    for i in range(sc.n_nodes):
        assert c.inspect_node(names[i])[1] == [int(x) for x in used[i]]
    assert (S.used_from_records(table, sc.n_nodes, sc.max_devs, ndev) == used).all()
