"""K1 (rebuild), K2 (events), K4 (bind) parity through the C ABI against the numpy restatement, which
tests/test_soa_reference.py checks against the object-level oracle."""
import numpy as np
import pytest

import gsx
from gsx import synth
from oracle import soa_reference as S
from tests import helpers as H

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n_nodes,n_devs,running,sort_by_node", [
    (1000, 8, 20000, False), (1000, 8, 20000, True), (5000, 8, 100000, False), (333, 5, 7000, True), (50, 16, 9000, False)])
def test_rebuild_matches_reference(n_nodes, n_devs, running, sort_by_node):
    sc = synth.make_cluster(n_nodes, n_devs, None, running, seed=n_nodes, quirks=True, fill=0.9)
    recs = sc.gs_pod_records()
    if sort_by_node:
        recs = recs[np.argsort(recs["node_idx"], kind="stable")]
    want = S.used_from_records(recs, n_nodes, n_devs, sc.dev_count())
    with gsx.Snapshot(n_nodes, n_devs, len(recs)) as s:
        s.node_upsert(sc.gs_nodes())
        s.pods_load(recs)
        s.rebuild()
        assert (s.read_used() == want).all()
        s.rebuild()                                               # idempotent
        assert (s.read_used() == want).all()
        back = s.read_pods(0, len(recs))
        legal = S.legal(recs, sc.dev_count())
        assert ((back["flags"] & 1).astype(bool) == (legal & ((recs["flags"] & 1) == 1))).all()


def test_rebuild_wraps_like_go_uint():
    with gsx.Snapshot(2, 2, 4) as s:
        s.upsert_nodes([0, 1], [100, 100], [2, 2], [0, 0])
        r = np.zeros(3, synth.POD_DT)
        r["mem"] = [2**63, 2**63, 5]
        r["node_idx"], r["dev"], r["flags"] = 0, 0, 3
        s.pods_load(r)
        s.rebuild()
        assert s.read_used()[0].tolist() == [5, 0]                # 2^63 + 2^63 wraps to 0
        bm = s.filter_batch([96, 95, 2**64 - 1])
        assert gsx.unpack_bitmap(bm, 2).tolist() == [[True, True], [True, True], [False, False]]


@pytest.mark.parametrize("seed", range(8))
def test_events_match_sequential_semantics_and_rebuild(seed):
    rng = np.random.default_rng(seed)
    n_nodes, n_devs, n_slots = 64, 4, 500
    sc = synth.make_cluster(n_nodes, n_devs, 16384, 0, seed=seed, quirks=True)
    ndev = sc.dev_count()
    table = np.zeros(n_slots, synth.POD_DT)
    used = np.zeros((n_nodes, n_devs), np.uint64)
    with gsx.Snapshot(n_nodes, n_devs, n_slots) as s:
        s.node_upsert(sc.gs_nodes())
        for batch in range(6):
            E = int(rng.integers(1, 900))
            ev = np.zeros(E, synth.EVENT_DT)
            ev["slot"] = rng.integers(0, n_slots, E)              # many duplicate slots per batch
            ev["op"] = rng.random(E) < 0.4
            ev["node_idx"] = rng.integers(0, n_nodes + 2, E)      # some beyond the snapshot
            ev["dev"] = rng.integers(-1, n_devs + 1, E)           # some illegal devices
            ev["mem"] = rng.choice(synth.CATALOGUE, E)
            ev["counted"] = rng.random(E) < 0.9
            S.apply_events(ev, table, used, ndev)
            s.pods_apply(ev)
            assert (s.read_used() == used).all(), (seed, batch)
            assert (s.read_pods(0, n_slots) == table).all()
        s.rebuild()                                               # K2(events) == K1(final table)
        assert (s.read_used() == used).all()


@pytest.mark.parametrize("hot", [False, True])
def test_bind_batch_matches_sequential_binpack(hot):
    sc = synth.make_config("C3")
    cap, used, ndev, cls, recs = H.soa_state(sc)
    rng = np.random.default_rng(3)
    B = 5000
    nodes = rng.integers(0, sc.n_nodes, B)
    if hot:                                                       # half the binds hit 1 % of the nodes
        hot_nodes = rng.integers(0, sc.n_nodes, sc.n_nodes // 100)
        pick = rng.random(B) < 0.5
        nodes[pick] = rng.choice(hot_nodes, int(pick.sum()))
    reqs = rng.choice(np.r_[synth.CATALOGUE, [0, -5, 1]], B).astype(np.int64)
    binds = synth.bind_records(nodes, reqs, first_slot=sc.n_pods, counted=rng.random(B) < 0.97)
    table = np.zeros(sc.n_pods + B, synth.POD_DT)
    table[: sc.n_pods] = recs
    table["flags"][: sc.n_pods][~S.legal(recs, ndev)] = 0
    want = S.bind_sequential(binds, cap, used, ndev, cls, table)
    with H.load_snapshot(sc, extra_pods=B) as s:
        got = s.bind_batch(binds)
        assert (got == want).all()
        assert (s.read_used() == used).all()
        newrecs = s.read_pods(sc.n_pods, B)
        ok = want >= 0
        assert (newrecs[ok] == table[sc.n_pods:][ok]).all() and (newrecs["flags"][~ok] == 0).all()
        # roll back every third successful bind: state returns to "never committed"
        undo = binds["slot"][ok][::3]
        S.apply_events(synth.remove_events(undo), table, used, ndev)
        s.bind_rollback(undo)
        assert (s.read_used() == used).all()
        s.rebuild()
        assert (s.read_used() == used).all()


def test_bind_calls_of_every_size_match_sequential_binpack():
    """gs_bind_batch takes three routes by queue length: one launch of one warp (<= 32 binds), the two kernels on the
    page-locked block in place (<= 1024), the two kernels on a device copy (more).  A long run of calls of mixed
    lengths, many binds per node (so queue order inside a node matters), must leave exactly what sequential binds do."""
    sc = synth.make_cluster(64, 8, 16384, running=300, seed=21, quirks=True, fill=0.6)
    cap, used, ndev, cls, recs = H.soa_state(sc)
    rng = np.random.default_rng(21)
    sizes = [1, 1, 2, 3, 31, 32, 33, 64, 200, 1, 1024, 1025, 32, 7, 1500, 1]
    total = sum(sizes)
    table = np.zeros(sc.n_pods + total, synth.POD_DT)
    table[: sc.n_pods] = recs
    table["flags"][: sc.n_pods][~S.legal(recs, ndev)] = 0
    with H.load_snapshot(sc, extra_pods=total) as s:
        slot = sc.n_pods
        for B in sizes:
            nodes = rng.integers(0, sc.n_nodes + 2, B)            # two names outside the snapshot: NODE_NOT_FOUND
            reqs = rng.choice(np.r_[synth.CATALOGUE[:6], [0, -5, 1, 16384]], B).astype(np.int64)
            binds = synth.bind_records(nodes, reqs, first_slot=slot, counted=rng.random(B) < 0.9)
            if B >= 3:
                binds["slot"][2] = sc.n_pods                       # a slot that already holds an entry (first call: fresh)
                if len(np.unique(binds["slot"])) != B:
                    binds["slot"][2] = slot + 2
            slot += B
            want = S.bind_sequential(binds, cap, used, ndev, cls, table)
            got = s.bind_batch(binds)
            assert (got == want).all(), B
            assert (s.read_used() == used).all(), B
        assert (s.read_pods(0, sc.n_pods + total) == table).all()
        s.rebuild()
        assert (s.read_used() == used).all()


def test_ka2_bind_through_the_abi():
    # docs/designs/designs.md:80
    with gsx.Snapshot(1, 4, 8) as s:
        s.upsert_nodes([0], [16276], [4], [0])
        s.pods_apply(synth.add_events([0, 1, 2], [0, 0, 0], [0, 1, 2], [4069, 8138, 12207]))
        assert s.bind_batch(synth.bind_records([0], [8138], first_slot=3)).tolist() == [1]
        assert s.read_used()[0].tolist() == [4069, 16276, 12207, 0]
        assert s.bind_batch(synth.bind_records([0, 0, 0, 5], [0, 16277, 16276, 1], first_slot=4)).tolist() == [-1, -1, 3, -2]


def test_inspect_and_node_rules():
    with gsx.Snapshot(4, 4, 4) as s:
        s.upsert_nodes([0, 1, 2], [33, 0, 10], [3, 4, 0], [0, 2, 0])      # KA-3-like truncation done by the host
        cap, used, cls = s.inspect(0)
        assert cap.tolist() == [33, 33, 33] and used.tolist() == [0, 0, 0] and cls == 0
        assert s.inspect(1)[2] == 2 and s.inspect(3)[2] == 1 and len(s.inspect(2)[0]) == 0
        bm = gsx.unpack_bitmap(s.filter_batch([0, 1, 34]), 4)
        assert bm.tolist() == [[True, False, False, False], [True, False, False, False], [False] * 4]
        with pytest.raises(gsx.GpushareError):
            s.upsert_nodes([4], [1], [1], [0])
        with pytest.raises(gsx.GpushareError):
            s.upsert_nodes([0], [1], [5], [0])
