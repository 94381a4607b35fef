"""K3 parity: gs_filter_batch (through the C ABI) against the CPU oracle.
Bit-exact: every word of the bitmap must equal the oracle's pass set."""
import os

import numpy as np
import pytest

import gsx
from gsx import synth
from oracle import oracle as O
from oracle import soa_reference as S
from tests import helpers as H

pytestmark = pytest.mark.gpu


def want_bitmap(sc, reqs, stride, used=None):
    cap, u, ndev, cls, _ = H.soa_state(sc)
    key, elig = S.node_keys(cap, u if used is None else used, ndev, cls)
    return S.pack_bitmap(S.filter_pass(np.asarray(reqs).astype(np.uint64), key, elig), stride)


def test_ka1_through_the_abi():
    # docs/designs/designs.md:64-68
    with gsx.Snapshot(3, 2, 8) as s:
        s.upsert_nodes([0, 1, 2], [16276] * 3, [2] * 3, [0] * 3)
        used = [(0, 0, 16276), (0, 1, 12207), (1, 0, 12207), (1, 1, 12207), (2, 0, 8138), (2, 1, 16176)]
        s.pods_apply(synth.add_events(range(6), [u[0] for u in used], [u[1] for u in used], [u[2] for u in used]))
        bm = s.filter_batch([8138])
        assert gsx.unpack_bitmap(bm, 3).tolist() == [[False, False, True]]
        assert s.read_used().tolist() == [[16276, 12207], [12207, 12207], [8138, 16176]]


def test_c1_filter_and_bind_against_object_oracle():
    """BASELINE config 1: 10 nodes x 4 GPUs x 16 GiB, 100 pods through Filter + Bind, every decision
    compared with the faithful object-level oracle (pass sets, FailedNodes classes, chosen GPU ids)."""
    sc = synth.make_cluster(10, 4, 16384, running=0, seed=0xC1, quirks=False)
    c = H.oracle_cluster(sc)
    names = H.all_names(sc) + ["ghost"]
    rng = np.random.default_rng(0xC1)
    reqs = rng.choice(np.r_[synth.CATALOGUE[:7], [0, 16385, 6000]], 100).astype(np.int64)
    with H.load_snapshot(sc, extra_pods=100) as s:
        cls = s.node_classes()
        for i, r in enumerate(reqs):
            pod = H.pending_pod(i, r)
            c.list_pod(pod)
            codes = c.filter_codes(pod, names)
            bits = gsx.unpack_bitmap(s.filter_batch([np.uint64(r)]), sc.n_nodes)[0]
            got = np.where(bits, O.PASS, O.INSUFFICIENT)
            got[cls == gsx.capi.NODE_NOT_SHARE] = O.NOT_SHARE
            got[cls == gsx.capi.NODE_NOT_FOUND] = O.NOT_FOUND
            assert (got == codes[:-1]).all() and codes[-1] == O.NOT_FOUND
            passing = np.flatnonzero(bits)
            target = int(passing[0]) if len(passing) else int(rng.integers(0, 10))
            dev_o, err = c.bind(f"pend-{i}", "default", f"uid-pend-{i}", names[target])
            dev_g = int(s.bind_batch(synth.bind_records([target], [r], first_slot=i))[0])
            assert dev_g == dev_o and (err == "") == (dev_g >= 0)
        for n in range(10):
            assert c.inspect_node(names[n])[1] == s.read_used(n, 1)[0].tolist()


def test_c2_against_object_oracle():
    """BASELINE config 2: 1k nodes x 8 devices (180 GiB), 20k running pods, 1k pending pods."""
    sc = synth.make_config("C2")
    c = H.oracle_cluster_tolerant(sc)
    names = H.all_names(sc)
    reqs = np.r_[synth.make_requests(500, "adversarial", 0xC2), synth.make_requests(500, "catalogue", 0xC2)]
    pods = [H.pending_pod(i, r) for i, r in enumerate(reqs)]
    codes = c.filter_many(pods, names, threads=8)
    with H.load_snapshot(sc) as s:
        for n in range(0, sc.n_nodes, 37):
            got = c.inspect_node(names[n])
            if got is not None:
                assert got[1] == s.read_used(n, 1)[0][: len(got[1])].tolist()
        bits = gsx.unpack_bitmap(s.filter_batch(reqs.astype(np.uint64)), sc.n_nodes)
        assert ((codes == O.PASS) == bits).all()
        cls = s.node_classes()
        assert ((codes == O.NOT_FOUND) == (cls == 1)[None, :]).all()
        assert ((codes == O.NOT_SHARE) == (cls == 2)[None, :]).all()


@pytest.mark.parametrize("seed", range(1, 41))
def test_fuzz_small_clusters(seed):
    rng = np.random.default_rng(seed)
    n_nodes = int(rng.integers(1, 65))
    n_devs = int(rng.choice([1, 2, 3, 4, 5, 8, 16]))
    sc = synth.make_cluster(n_nodes, n_devs, 16384, running=int(rng.integers(0, 6 * n_nodes * n_devs // 2 + 1)),
                            seed=seed, quirks=True, fill=float(rng.uniform(0.3, 1.1)))
    P = int(rng.integers(1, 200))
    reqs = synth.make_requests(P, "adversarial", seed, max_req=20000)
    reqs[rng.random(P) < 0.05] = -1
    reqs[rng.random(P) < 0.05] = 2**62
    stride = (n_nodes + 31) // 32 + int(rng.integers(0, 6))
    with H.load_snapshot(sc) as s:
        got = s.filter_batch(reqs.astype(np.uint64), row_stride_words=stride)
        want = want_bitmap(sc, reqs.astype(np.uint64), stride)
        w = (n_nodes + 31) // 32
        assert (got[:, :w] == want[:, :w]).all()
        assert (got[:, w:] == 0).all()
    if seed <= 10:  # and straight against the object-level oracle
        c = H.oracle_cluster_tolerant(sc)
        names = H.all_names(sc)
        bits = gsx.unpack_bitmap(got, n_nodes)
        for p in range(min(P, 25)):
            assert ((c.filter_codes(H.pending_pod(p, reqs[p]), names) == O.PASS) == bits[p]).all()


def test_edge_shapes():
    sc = synth.make_cluster(1000, 8, 184320, running=4000, seed=9, quirks=True)
    with H.load_snapshot(sc) as s:
        assert s.filter_batch(np.zeros(0, np.uint64)).shape[0] == 0              # empty batch
        w = s.bitmap_words
        for P, stride in [(1, w), (1, w + 1), (63, 32), (64, 36), (65, 33), (129, 35)]:
            reqs = synth.make_requests(P, "adversarial", P)
            got = s.filter_batch(reqs.astype(np.uint64), row_stride_words=stride)
            assert (got[:, :w] == want_bitmap(sc, reqs, stride)[:, :w]).all(), (P, stride)
        same = np.full(300, 4096, np.uint64)                                     # one distinct row
        assert (s.filter_batch(same)[:, :w] == want_bitmap(sc, same, (w + 3) // 4 * 4)[:, :w]).all()
        with pytest.raises(gsx.GpushareError):
            s.filter_batch(np.ones(4, np.uint64), row_stride_words=w - 1)


def test_batches_larger_than_one_launch():
    sc = synth.make_cluster(300, 8, 184320, running=3000, seed=11, quirks=True, fill=0.8)
    reqs = synth.make_requests(2 * 16384 + 1234, "adversarial", 3).astype(np.uint64)
    with H.load_snapshot(sc) as s:
        got = s.filter_batch(reqs)
        assert (got == want_bitmap(sc, reqs, got.shape[1])).all()


def test_full_u64_range_keys_sort_correctly():
    sc = synth.make_cluster(200, 8, 184320, running=1500, seed=12, quirks=True, fill=1.2)
    rng = np.random.default_rng(5)
    reqs = rng.integers(0, 2**64 - 1, 3000, dtype=np.uint64)
    reqs[::7] = rng.integers(0, 200000, len(reqs[::7]), dtype=np.uint64)
    reqs[::11] = np.uint64(2**64 - 1)
    with H.load_snapshot(sc) as s:
        got = s.filter_batch(reqs)
        assert (got == want_bitmap(sc, reqs, got.shape[1])).all()


@pytest.mark.parametrize("cfg,kind", [("C3", "adversarial"), ("C3", "catalogue"), ("C4", "adversarial"), ("C4", "catalogue")])
def test_full_size_configs(cfg, kind):
    """BASELINE configs 3 and 4 at full size against the numpy reference, plus size-independent
    properties: monotonicity in the request, req = 0 row == eligibility mask, popcount checksum."""
    sc = synth.make_config(cfg)
    P = synth.CONFIGS[cfg]["pending"]
    reqs = synth.make_requests(P, kind, synth.CONFIGS[cfg]["seed"]).astype(np.uint64)
    cap, used, ndev, cls, _ = H.soa_state(sc)
    key, elig = S.node_keys(cap, used, ndev, cls)
    with H.load_snapshot(sc) as s:
        assert (s.read_used() == used).all()
        got = s.filter_batch(reqs)
        stride = got.shape[1]
        for p0 in range(0, P, 1000):
            want = S.pack_bitmap(S.filter_pass(reqs[p0:p0 + 1000], key, elig), stride)
            assert (got[p0:p0 + 1000] == want).all(), (cfg, kind, p0)
        order = np.argsort(reqs, kind="stable")
        rows = got[order]
        assert ((rows[1:] & ~rows[:-1]) == 0).all()                       # larger request => subset
        zero = np.flatnonzero(reqs == 0)
        if len(zero):
            assert (got[zero[0]] == S.pack_bitmap(elig[None, :], stride)[0]).all()


def test_pipelined_back_to_back_filters_are_identical():
    """GS_OPT_PIPELINE_FILTERS lets the preparation kernel of one batch run under the tail of the previous batch's
    bitmap kernel.  Twelve different batches are enqueued back to back (no host synchronisation in between) into
    rotating output buffers, plus one call that reuses the previous output (which must fall back to stream order);
    every bitmap must equal the reference."""
    import torch

    sc = synth.make_config("C3")
    cap, used, ndev, cls, _ = H.soa_state(sc)
    key, elig = S.node_keys(cap, used, ndev, cls)
    P = 3000
    with H.load_snapshot(sc) as s:
        s.set_option(gsx.capi.OPT_PIPELINE_FILTERS, 1)
        stride = (s.bitmap_words + 3) // 4 * 4
        st = torch.cuda.Stream()
        batches = [synth.make_requests(P, "adversarial" if i % 2 else "catalogue", 100 + i).astype(np.uint64) for i in range(12)]
        d_reqs = [torch.from_numpy(b.view(np.int64)).cuda() for b in batches]
        outs = [torch.zeros((P, stride), dtype=torch.int32, device="cuda") for _ in range(3)]
        torch.cuda.synchronize()
        got = []
        for i, d_req in enumerate(d_reqs):
            o = outs[i % 3] if i != 7 else outs[(i - 1) % 3]        # call 7 overwrites call 6's buffer
            s.filter_batch_device(d_req.data_ptr(), P, o.data_ptr(), stride, st.cuda_stream)
            if i % 3 == 2 or i in (6, 7):                            # read back before the buffer is reused
                st.synchronize()
                for j in ([i] if i in (6, 7) else [i - 2, i - 1, i]):
                    if j not in (6, 7) or j == i:
                        src = o if j == i else outs[j % 3]
                        got.append((j, src.cpu().numpy().view(np.uint32).copy()))
        st.synchronize()
        seen = set()
        for j, bm in got:
            want = S.pack_bitmap(S.filter_pass(batches[j], key, elig), stride)
            assert (bm == want).all(), j
            seen.add(j)
        assert seen == set(range(12))


def test_pipelined_mixed_sizes_and_state_changes():
    """Overlapped batches of very different sizes (grids from one CTA to several waves), each with its own output
    buffer, enqueued back to back; half way through the first round, pod events change the snapshot (which ends
    the chain of overlapping filters) and the batches after it must see the new state.  Every bitmap is compared
    with the reference afterwards."""
    import torch

    sc = synth.make_config("C3")
    cap, used, ndev, cls, _ = H.soa_state(sc)
    key0, elig0 = S.node_keys(cap, used, ndev, cls)
    sizes = [1, 64, 3000, 7, 2048, 65, 12288, 130, 1, 5000, 640, 3]
    rng = np.random.default_rng(7)
    with H.load_snapshot(sc, extra_pods=4096) as s:
        s.set_option(gsx.capi.OPT_PIPELINE_FILTERS, 1)
        stride = (s.bitmap_words + 7) // 8 * 8
        st = torch.cuda.Stream()
        batches = [synth.make_requests(n, "adversarial" if i % 3 else "catalogue", 300 + i).astype(np.uint64)
                   for i, n in enumerate(sizes)]
        d_reqs = [torch.from_numpy(b.view(np.int64)).cuda() for b in batches]
        outs = [torch.full((n, stride), -1, dtype=torch.int32, device="cuda") for n in sizes]
        # events applied between batch 5 and batch 6: fill one device of many nodes
        n_ev = 2000
        nodes = rng.choice(sc.n_nodes, n_ev, replace=False).astype(np.uint32)
        ev = synth.add_events(sc.n_pods + np.arange(n_ev, dtype=np.uint32), nodes, np.zeros(n_ev, np.int16),
                              np.full(n_ev, 150000, np.uint64))
        torch.cuda.synchronize()
        for rep in range(3):  # the same sequence three times: epochs and buffer sets keep alternating
            for i, d_req in enumerate(d_reqs):
                if rep == 0 and i == 6:
                    st.synchronize()   # calls on one snapshot are ordered by the caller: the filters above read `used`
                    s.pods_apply(ev)   # (synchronous on the snapshot's own stream)
                s.filter_batch_device(d_req.data_ptr(), sizes[i], outs[i].data_ptr(), stride, st.cuda_stream)
            st.synchronize()
            if rep == 0:
                first = [o.cpu().numpy().view(np.uint32).copy() for o in outs]
        last = [o.cpu().numpy().view(np.uint32).copy() for o in outs]
        used1 = s.read_used().reshape(used.shape)
        key1, elig1 = S.node_keys(cap, used1, ndev, cls)
        assert (key1 != key0).any()
        for i in range(len(sizes)):
            before = S.pack_bitmap(S.filter_pass(batches[i], key0, elig0), stride)
            after = S.pack_bitmap(S.filter_pass(batches[i], key1, elig1), stride)
            assert (first[i] == (before if i < 6 else after)).all(), ("first pass", i)
            assert (last[i] == after).all(), ("last pass", i)


def test_pipelined_c4_hard_request_distributions():
    """Full-size config C4 with the request distributions that stress the tile classification: every row inside
    the crowded key buckets (all tiles dense), every row above the largest capacity (open-ended last bucket),
    requests that do not fit 32 bits (wrapped negative sums: the 64-bit key path), zeros, and a mix.  Ten thousand
    pods each, enqueued back to back with pipelining on; bit-exact against the flat-array restatement."""
    import torch

    sc = synth.make_config("C4")
    cap, used, ndev, cls, _ = H.soa_state(sc)
    key, elig = S.node_keys(cap, used, ndev, cls)
    P = 10000
    rng = np.random.default_rng(11)
    wide = rng.integers(1, 184320, P).astype(np.uint64)
    wide[::7] = np.uint64(2**64 - 1) - rng.integers(0, 5000, len(wide[::7])).astype(np.uint64)  # uint(negative sum)
    wide[3::11] = np.uint64(2**40) + rng.integers(0, 99, len(wide[3::11])).astype(np.uint64)
    mix = np.concatenate([rng.integers(180000, 184321, P // 2), rng.integers(0, 3, P // 4), rng.integers(150000, 400000, P - P // 2 - P // 4)]).astype(np.uint64)
    rng.shuffle(mix)
    batches = [
        rng.integers(180000, 184321, P).astype(np.uint64),   # all rows in the crowded buckets
        np.full(P, 1_000_000, dtype=np.uint64),               # open-ended bucket only
        wide,
        np.zeros(P, dtype=np.uint64),
        mix,
    ]
    with H.load_snapshot(sc) as s:
        s.set_option(gsx.capi.OPT_PIPELINE_FILTERS, 1)
        stride = (s.bitmap_words + 7) // 8 * 8
        st = torch.cuda.Stream()
        d_reqs = [torch.from_numpy(b.view(np.int64)).cuda() for b in batches]
        outs = [torch.full((P, stride), -1, dtype=torch.int32, device="cuda") for _ in batches]
        torch.cuda.synchronize()
        for rep in range(2):
            for d_req, o in zip(d_reqs, outs):
                s.filter_batch_device(d_req.data_ptr(), P, o.data_ptr(), stride, st.cuda_stream)
        st.synchronize()
        for i, (b, o) in enumerate(zip(batches, outs)):
            got = o.cpu().numpy().view(np.uint32)
            want = S.pack_bitmap(S.filter_pass(b, key, elig), stride)
            assert (got == want).all(), i
        # Overlapping batches: the expansion kernel's CTAs take their batch's preparation to be complete from the launch
        # order and confirm it by an epoch; none may ever find it unreached.  (A build with -DGS_TEST_FORCE_EARLY_START
        # pretends an eighth of them do: the same bitmaps must come out, through the wait-and-ask-again path.)
        forced = os.environ.get("GS_EXPECT_EARLY_STARTS", "0") == "1"
        assert (s.early_starts > 0) if forced else (s.early_starts == 0)
