"""The two other shapes of the Filter answer, through the C ABI, bit-exact:

* gs_filter_batch_compact: rank[n] / cutoff[p]; expanding them on the host (rank >= cutoff) must reproduce the
  bitmap of gs_filter_batch bit for bit, and both must equal the CPU restatement;
* gs_filter_batch_subset: the candidate list of one Filter request (pkg/scheduler/predicate.go:21).
"""
import numpy as np
import pytest

import gsx
from gsx import synth
from oracle import soa_reference as S
from tests import helpers as H

pytestmark = pytest.mark.gpu


def hard_requests(P, rng):
    """Every request class the path distinguishes: catalogue sizes, uniform values, zeros, values above every
    capacity, wrapped negative sums (>= 2^63), values just above 2^32, heavy duplicates."""
    r = rng.integers(1, 184321, P).astype(np.uint64)
    r[::5] = rng.choice(synth.CATALOGUE, len(r[::5])).astype(np.uint64)
    r[1::13] = 0
    r[2::17] = np.uint64(2**64 - 1) - rng.integers(0, 5000, len(r[2::17])).astype(np.uint64)
    r[3::19] = np.uint64(2**32) + rng.integers(0, 99, len(r[3::19])).astype(np.uint64)
    r[4::23] = 1_000_000
    r[5::7] = 8192
    return r


@pytest.mark.parametrize("cfg,kind", [("C2", "adversarial"), ("C2", "hard"), ("C3", "adversarial"), ("C3", "catalogue"),
                                      ("C3", "hard"), ("C4", "adversarial"), ("C4", "catalogue"), ("C4", "hard")])
def test_compact_expands_to_the_bitmap(cfg, kind):
    sc = synth.make_config(cfg)
    P = synth.CONFIGS[cfg]["pending"]
    rng = np.random.default_rng(synth.CONFIGS[cfg]["seed"])
    reqs = hard_requests(P, rng) if kind == "hard" else synth.make_requests(P, kind, synth.CONFIGS[cfg]["seed"]).astype(np.uint64)
    cap, used, ndev, cls, _ = H.soa_state(sc)
    key, elig = S.node_keys(cap, used, ndev, cls)
    with H.load_snapshot(sc) as s:
        bitmap = s.filter_batch(reqs)
        rank, cutoff = s.filter_batch_compact(reqs)
        assert rank.shape == (1, sc.n_nodes) and cutoff.shape == (P,)
        expanded = gsx.expand_compact(rank, cutoff, sc.n_nodes, bitmap.shape[1])
        assert (expanded == bitmap).all(), "host expansion of the compact form differs from gs_filter_batch"
        for p0 in range(0, P, 1000):
            want = S.pack_bitmap(S.filter_pass(reqs[p0:p0 + 1000], key, elig), bitmap.shape[1])
            assert (expanded[p0:p0 + 1000] == want).all(), (cfg, kind, p0)
        # what the two arrays are, not only what they imply
        srt = np.sort(reqs)
        assert (cutoff == 1 + np.searchsorted(srt, reqs, side="left")).all()
        assert (rank[0] == np.where(elig, np.searchsorted(srt, key, side="right"), 0)).all()


def test_compact_edge_shapes():
    sc = synth.make_cluster(333, 4, 16384, running=1500, seed=4, quirks=True, fill=0.9)
    cap, used, ndev, cls, _ = H.soa_state(sc)
    key, elig = S.node_keys(cap, used, ndev, cls)
    rng = np.random.default_rng(4)
    with H.load_snapshot(sc) as s:
        # one pod; all pods equal; a batch longer than one launch handles (ranks per chunk); padded rank rows
        for reqs, stride in [(np.array([4096], np.uint64), None), (np.full(5000, 2048, np.uint64), None),
                             (hard_requests(gsx.capi.MAX_BATCH + 3000, rng), 400), (np.zeros(77, np.uint64), 336)]:
            rank, cutoff = s.filter_batch_compact(reqs, rank_stride=stride)
            chunks = (len(reqs) + gsx.capi.MAX_BATCH - 1) // gsx.capi.MAX_BATCH
            assert rank.shape == (chunks, stride or sc.n_nodes)
            assert (rank[:, sc.n_nodes:] == 0).all()
            got = gsx.unpack_bitmap(gsx.expand_compact(rank, cutoff, sc.n_nodes), sc.n_nodes)
            assert (got == S.filter_pass(reqs, key, elig)).all()
        rank, cutoff = s.filter_batch_compact(np.zeros(0, np.uint64))
        assert len(cutoff) == 0
        with pytest.raises(gsx.GpushareError):
            s.filter_batch_compact(np.ones(3, np.uint64), rank_stride=sc.n_nodes - 1)


def test_compact_device_variant_and_state_changes():
    import torch

    sc = synth.make_config("C3")
    cap, used, ndev, cls, recs = H.soa_state(sc)
    P = 4000
    reqs = synth.make_requests(P, "adversarial", 9).astype(np.uint64)
    with H.load_snapshot(sc, extra_pods=4096) as s:
        st = torch.cuda.Stream()
        d_req = torch.from_numpy(reqs.view(np.int64)).cuda()
        d_rank = torch.zeros(sc.n_nodes, dtype=torch.int16, device="cuda")
        d_cut = torch.zeros(P, dtype=torch.int16, device="cuda")
        torch.cuda.synchronize()
        table = np.zeros(sc.n_pods + 4096, synth.POD_DT)
        table[: sc.n_pods] = recs
        table["flags"][: sc.n_pods][~S.legal(recs, ndev)] = 0
        rng = np.random.default_rng(2)
        for round_ in range(4):
            if round_:
                # events queued asynchronously on the snapshot's stream; the filter on another stream must see them
                ev = synth.add_events(sc.n_pods + rng.integers(0, 4096, 500), rng.integers(0, sc.n_nodes, 500),
                                      rng.integers(0, 8, 500), rng.choice(synth.CATALOGUE[3:8], 500))
                S.apply_events(ev, table, used, ndev)
                s.pods_apply_async(ev)
            s.filter_batch_compact_device(d_req.data_ptr(), P, d_rank.data_ptr(), sc.n_nodes, d_cut.data_ptr(), st.cuda_stream)
            st.synchronize()
            key, elig = S.node_keys(cap, used, ndev, cls)
            rank = d_rank.cpu().numpy().view(np.uint16)
            cut = d_cut.cpu().numpy().view(np.uint16)
            assert ((rank[None, :] >= cut[:, None]) == S.filter_pass(reqs, key, elig)).all(), round_
        assert (s.read_used() == used).all()


@pytest.mark.parametrize("seed", range(4))
def test_subset_matches_reference(seed):
    rng = np.random.default_rng(seed)
    sc = synth.make_cluster(1500, 8, 184320, running=9000, seed=40 + seed, quirks=True, fill=0.9)
    cap, used, ndev, cls, _ = H.soa_state(sc)
    key, elig = S.node_keys(cap, used, ndev, cls)
    full = None
    with H.load_snapshot(sc) as s:
        for P, C in [(1, 30), (1, 1), (7, 33), (300, 1500), (2000, 257), (3, 4000)]:
            reqs = hard_requests(max(P, 40), rng)[:P]
            sub = rng.integers(0, sc.n_nodes, C).astype(np.uint32)      # duplicates allowed, any order
            if C > 20:
                sub[3] = sc.n_nodes + 5                                  # names no node of the snapshot
                sub[7] = 2**32 - 1
            stride = (C + 31) // 32 + (seed % 2)                         # with and without padding words
            got = s.filter_batch_subset(reqs, sub, row_stride_words=stride)
            ok = sub < sc.n_nodes
            safe = np.where(ok, sub, 0)
            want = (key[safe][None, :] >= reqs[:, None]) & (elig[safe] & ok)[None, :]
            assert (got == S.pack_bitmap(want, stride)).all(), (P, C)
        # the identity subset is the full filter
        reqs = hard_requests(500, rng)
        full = s.filter_batch(reqs)
        ident = s.filter_batch_subset(reqs, np.arange(sc.n_nodes, dtype=np.uint32))
        assert (ident == full[:, : ident.shape[1]]).all()
        assert s.filter_batch_subset(reqs, np.zeros(0, np.uint32)).sum() == 0
        with pytest.raises(gsx.GpushareError):
            s.filter_batch_subset(reqs, np.arange(100, dtype=np.uint32), row_stride_words=3)


def test_subset_at_c4_scale():
    """One Filter request as kube-scheduler sends it at config C4's size: a few thousand candidate names out of
    50k nodes, one pod; and a batch of pods over the same candidates."""
    sc = synth.make_config("C4")
    cap, used, ndev, cls, _ = H.soa_state(sc)
    key, elig = S.node_keys(cap, used, ndev, cls)
    rng = np.random.default_rng(8)
    with H.load_snapshot(sc) as s:
        sub = rng.choice(sc.n_nodes, 5000, replace=False).astype(np.uint32)
        for P in (1, 1000):
            reqs = synth.make_requests(P, "adversarial", 77).astype(np.uint64)
            got = s.filter_batch_subset(reqs, sub)
            want = (key[sub][None, :] >= reqs[:, None]) & elig[sub][None, :]
            assert (got == S.pack_bitmap(want, got.shape[1])).all()


def test_compact_gather_entry_point_writes_every_destination():
    """gs_filter_batch_compact_device_gather on one GPU: two destination arrays stand in for "this GPU's copy and a peer's
    copy" of the gathered [shards][slice] array; the shard's ranks must land at the given offset of both, zero padded
    to the slice length, and nothing else may be touched."""
    import torch

    sc = synth.make_cluster(1000, 8, 184320, running=6000, seed=6, quirks=True, fill=0.9)
    cap, used, ndev, cls, _ = H.soa_state(sc)
    key, elig = S.node_keys(cap, used, ndev, cls)
    P, slice_len, offset = 3000, 1024, 1024                      # shard 1 of 3
    reqs = hard_requests(P, np.random.default_rng(6))
    with H.load_snapshot(sc) as s:
        d_req = torch.from_numpy(reqs.view(np.int64)).cuda()
        dsts = [torch.full((3 * slice_len,), -1, dtype=torch.int16, device="cuda") for _ in range(2)]
        d_cut = torch.zeros(P, dtype=torch.int16, device="cuda")
        st = torch.cuda.Stream()
        torch.cuda.synchronize()
        s.filter_batch_compact_device_gather(d_req.data_ptr(), P, [d.data_ptr() for d in dsts], offset, slice_len,
                                             d_cut.data_ptr(), st.cuda_stream)
        st.synchronize()
        srt = np.sort(reqs)
        want = np.full(3 * slice_len, 0xFFFF, np.uint16)
        want[offset: offset + slice_len] = 0
        want[offset: offset + sc.n_nodes] = np.where(elig, np.searchsorted(srt, key, side="right"), 0)
        for d in dsts:
            assert (d.cpu().numpy().view(np.uint16) == want).all()
        assert (d_cut.cpu().numpy().view(np.uint16) == 1 + np.searchsorted(srt, reqs, side="left")).all()
        with pytest.raises(gsx.GpushareError):
            s.filter_batch_compact_device_gather(d_req.data_ptr(), P, [dsts[0].data_ptr()], 0, sc.n_nodes - 1, d_cut.data_ptr(), 0)


def test_compact_with_page_locked_buffers_is_the_same_answer():
    """With page-locked caller buffers the kernels read the requests and write rank[] / cutoff[] in the caller's memory
    directly (no staging copies); the answer must be the one the staged path gives, for one chunk, for several chunks,
    for a sub-array of a larger page-locked block, after a state change, and when only some of the three buffers are
    page-locked (staged path again)."""
    sc = synth.make_config("C3")
    cap, used, ndev, cls, recs = H.soa_state(sc)
    rng = np.random.default_rng(12)
    N = sc.n_nodes
    with H.load_snapshot(sc, extra_pods=2048) as s:
        for P in (1, 777, gsx.capi.MAX_BATCH, gsx.capi.MAX_BATCH + 5000):
            reqs = hard_requests(max(P, 40), rng)[:P]
            chunks = (P + gsx.capi.MAX_BATCH - 1) // gsx.capi.MAX_BATCH
            want_rank, want_cut = s.filter_batch_compact(reqs)                     # pageable numpy arrays: staged
            h_req = gsx.PinnedBuffer(P + 3, np.uint64)
            h_rank = gsx.PinnedBuffer((chunks + 1, N), np.uint16)
            h_cut = gsx.PinnedBuffer(P + 8, np.uint16)
            h_req.array[3:] = reqs
            h_rank.array[:] = 0xABCD
            h_cut.array[:] = 0xABCD
            rank, cut = s.filter_batch_compact(h_req.array[3:], out_rank=h_rank.array[1:], out_cutoff=h_cut.array[8:])
            assert (rank == want_rank).all() and (cut == want_cut).all(), P
            assert (h_rank.array[0] == 0xABCD).all() and (h_cut.array[:8] == 0xABCD).all()   # nothing else touched
            # mixed: only the requests page-locked
            rank2, cut2 = s.filter_batch_compact(h_req.array[3:])
            assert (rank2 == want_rank).all() and (cut2 == want_cut).all()
            for b in (h_req, h_rank, h_cut):
                b.free()
        # a write between two page-locked calls is seen by the second
        P = 3000
        reqs = hard_requests(P, rng)
        h_req = gsx.PinnedBuffer(P, np.uint64)
        h_rank = gsx.PinnedBuffer((1, N), np.uint16)
        h_cut = gsx.PinnedBuffer(P, np.uint16)
        h_req.array[:] = reqs
        s.filter_batch_compact(h_req.array, out_rank=h_rank.array, out_cutoff=h_cut.array)
        table = np.zeros(sc.n_pods + 2048, synth.POD_DT)
        table[: sc.n_pods] = recs
        table["flags"][: sc.n_pods][~S.legal(recs, ndev)] = 0
        ev = synth.add_events(sc.n_pods + np.arange(600), rng.integers(0, N, 600), rng.integers(0, 8, 600),
                              rng.choice(synth.CATALOGUE[3:8], 600))
        S.apply_events(ev, table, used, ndev)
        s.pods_apply_async(ev)
        rank, cut = s.filter_batch_compact(h_req.array, out_rank=h_rank.array, out_cutoff=h_cut.array)
        key, elig = S.node_keys(cap, used, ndev, cls)
        assert ((rank[0][None, :] >= cut[:, None]) == S.filter_pass(reqs, key, elig)).all()
