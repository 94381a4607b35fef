"""Differential fuzz of a snapshot's whole life on small clusters ("randomized differential testing
oracle <-> GPU ... at N <= 64"): random interleavings of pod-event batches (waited for and queued), bind batches of every
length class (one-warp, in place, staged), bind rollbacks, rebuilds, and Filters in all three forms of the answer --
bitmap, candidate list, rank/cutoff -- with pageable and page-locked buffers, against the flat-array restatement
(oracle/soa_reference.py) stepped through the same operations.  Everything bit for bit."""
import numpy as np
import pytest

import gsx
from gsx import synth
from oracle import soa_reference as S
from tests import helpers as H

pytestmark = pytest.mark.gpu


def _requests(rng, P, cap_max):
    r = rng.integers(0, 2 * cap_max, P).astype(np.uint64)
    r[rng.random(P) < 0.1] = rng.choice(synth.CATALOGUE, int((rng.random(P) < 0.1).sum()) or 1)[0]
    r[rng.random(P) < 0.03] = 0
    r[rng.random(P) < 0.03] = np.uint64(2**64 - 1) - np.uint64(rng.integers(0, 100))   # uint(negative sum)
    r[rng.random(P) < 0.03] = np.uint64(2**33) + np.uint64(rng.integers(0, 100))
    return r


@pytest.mark.parametrize("seed", range(1, 49))
def test_random_life_of_a_small_snapshot(seed):
    rng = np.random.default_rng(1000 + seed)
    n_nodes = int(rng.integers(1, 65))
    n_devs = int(rng.choice([1, 2, 4, 8]))
    cap0 = int(rng.choice([4096, 16384, 184320]))
    sc = synth.make_cluster(n_nodes, n_devs, cap0, running=int(rng.integers(0, 3 * n_nodes * n_devs + 1)), seed=seed, quirks=True,
                            fill=float(rng.uniform(0.3, 1.1)))
    cap, used, ndev, cls, recs = H.soa_state(sc)
    extra = 6000
    table = np.zeros(sc.n_pods + extra, synth.POD_DT)
    table[: sc.n_pods] = recs
    table["flags"][: sc.n_pods][~S.legal(recs, ndev)] = 0
    next_slot = sc.n_pods
    cap_max = int(cap.max()) if len(cap) else 1
    pinned = []
    with H.load_snapshot(sc, extra_pods=extra) as s:
        for step in range(int(rng.integers(8, 20))):
            op = rng.choice(["events", "events_async", "bind", "rollback", "rebuild", "filter", "subset", "compact", "compact_pinned"],
                            p=[0.12, 0.12, 0.2, 0.06, 0.05, 0.12, 0.13, 0.1, 0.1])
            if op in ("events", "events_async"):
                E = int(rng.integers(1, 120))
                adds = rng.random(E) < 0.6
                slots = np.where(adds, sc.n_pods + rng.integers(0, extra, E), rng.integers(0, max(1, next_slot), E))
                ev = synth.add_events(slots, rng.integers(0, n_nodes, E), rng.integers(-1, n_devs + 1, E),
                                      rng.choice(synth.CATALOGUE[:7], E), counted=rng.random(E) < 0.9)
                ev["op"][~adds] = gsx.capi.EV_REMOVE
                S.apply_events(ev, table, used, ndev)
                (s.pods_apply if op == "events" else s.pods_apply_async)(ev)
            elif op == "bind":
                B = int(rng.choice([1, 1, 2, 5, 31, 32, 33, 100, 1025]))
                if next_slot + B > sc.n_pods + extra:
                    continue
                nodes = rng.integers(0, n_nodes + 1, B)                      # one name outside the snapshot
                reqs = rng.choice(np.r_[synth.CATALOGUE[:7], [0, -3, 1, cap0]], B).astype(np.int64)
                binds = synth.bind_records(nodes, reqs, first_slot=next_slot, counted=rng.random(B) < 0.9)
                next_slot += B
                want = S.bind_sequential(binds, cap, used, ndev, cls, table)
                got = s.bind_batch(binds)
                assert (got == want).all(), (seed, step, B)
            elif op == "rollback":
                k = int(rng.integers(1, 20))
                slots = np.unique(rng.integers(sc.n_pods, max(sc.n_pods + 1, next_slot), k)).astype(np.uint32)
                S.apply_events(synth.remove_events(slots), table, used, ndev)
                s.bind_rollback(slots)
            elif op == "rebuild":
                s.rebuild()
            else:
                key, elig = S.node_keys(cap, used, ndev, cls)
                P = int(rng.choice([1, 1, 3, 64, 65, 300]))
                reqs = _requests(rng, P, cap_max)
                if op == "filter":
                    stride = (n_nodes + 31) // 32 + int(rng.integers(0, 3))
                    got = s.filter_batch(reqs, row_stride_words=stride)
                    assert (got == S.pack_bitmap(S.filter_pass(reqs, key, elig), stride)).all(), (seed, step)
                elif op == "subset":
                    C = int(rng.choice([1, 5, 30, 64, 200, 3000]))
                    sub = rng.integers(0, n_nodes + 2, C).astype(np.uint32)     # two indices name no node
                    stride = (C + 31) // 32 + int(rng.integers(0, 2))
                    got = s.filter_batch_subset(reqs, sub, row_stride_words=stride)
                    ok = sub < n_nodes
                    safe = np.where(ok, sub, 0)
                    want = (key[safe][None, :] >= reqs[:, None]) & (elig[safe] & ok)[None, :]
                    assert (got == S.pack_bitmap(want, stride)).all(), (seed, step, P, C)
                else:
                    if op == "compact_pinned":
                        h_req = gsx.PinnedBuffer(P, np.uint64)
                        h_rank = gsx.PinnedBuffer((1, n_nodes), np.uint16)
                        h_cut = gsx.PinnedBuffer(P, np.uint16)
                        pinned += [h_req, h_rank, h_cut]
                        h_req.array[:] = reqs
                        rank, cut = s.filter_batch_compact(h_req.array, out_rank=h_rank.array, out_cutoff=h_cut.array)
                    else:
                        rank, cut = s.filter_batch_compact(reqs)
                    assert ((rank[0][None, :] >= cut[:, None]) == S.filter_pass(reqs, key, elig)).all(), (seed, step)
        assert (s.read_used() == used).all()
        assert (s.read_pods(0, len(table)) == table).all()
        s.rebuild()
        assert (s.read_used() == used).all()
        assert s.early_starts == 0
    for b in pinned:
        b.free()
