"""The boundary under concurrency (SURVEY.md 8b "Threading"): Filters from several host threads, Filters beside
writers, two snapshots sharing one GPU, filters beside a foreign kernel that fills the GPU, and the slot rules of
gs_bind_batch.  Everything through the C ABI; results bit-exact against the CPU restatement."""
import threading

import numpy as np
import pytest

import gsx
from gsx import ext as X
from gsx import synth
from oracle import oracle as O
from oracle import soa_reference as S
from tests import helpers as H

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


def _run_threads(fns):
    errs = []

    def wrap(f):
        def g():
            try:
                f()
            except BaseException as e:  # noqa: BLE001
                errs.append(e)
        return g

    ts = [threading.Thread(target=wrap(f)) for f in fns]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    if errs:
        raise errs[0]


def test_host_filters_from_several_threads():
    """gs_filter_batch / _compact / _subset from six threads at once on one snapshot: each call runs in its own
    filter context (stream, staging, work buffers), as the reference's Filters share a node under RLock."""
    sc = synth.make_config("C3")
    cap, used, ndev, cls, _ = H.soa_state(sc)
    key, elig = S.node_keys(cap, used, ndev, cls)
    with H.load_snapshot(sc) as s:
        def worker(t):
            def run():
                rng = np.random.default_rng(t)
                for it in range(6):
                    reqs = synth.make_requests(700 + 100 * t, "adversarial" if (t + it) % 2 else "catalogue", 50 * t + it).astype(np.uint64)
                    want = S.filter_pass(reqs, key, elig)
                    if it % 3 == 0:
                        got = s.filter_batch(reqs)
                        assert (got == S.pack_bitmap(want, got.shape[1])).all(), (t, it)
                    elif it % 3 == 1:
                        rank, cut = s.filter_batch_compact(reqs)
                        assert ((rank[0][None, :] >= cut[:, None]) == want).all(), (t, it)
                    else:
                        sub = rng.choice(sc.n_nodes, 900, replace=False).astype(np.uint32)
                        got = s.filter_batch_subset(reqs, sub)
                        assert (got == S.pack_bitmap(want[:, sub], got.shape[1])).all(), (t, it)
            return run

        _run_threads([worker(t) for t in range(6)])


def test_filters_beside_event_batches_see_a_prefix_of_the_writes():
    """One thread applies event batches (each fills more nodes), three threads filter all the while.  Writers are
    exclusive, Filters share: every bitmap must be the answer for the state after some whole number of batches."""
    sc = synth.make_cluster(2000, 8, 184320, running=0, seed=3, quirks=False)
    cap, ndev, cls = sc.dev_capacity(), sc.dev_count(), sc.node_class()
    reqs = np.array([1, 40000, 100000, 184320], np.uint64)
    K = 12
    used = np.zeros((sc.n_nodes, 8), np.uint64)
    batches, states = [], []
    key, elig = S.node_keys(cap, used, ndev, cls)
    states.append(S.filter_pass(reqs, key, elig))
    for k in range(K):
        nodes = np.arange(k * 150, (k + 1) * 150)
        ev = synth.add_events(np.arange(k * 1200, (k + 1) * 1200), np.repeat(nodes, 8), np.tile(np.arange(8), 150),
                              np.full(1200, 90000 + 1000 * k))
        batches.append(ev)
        used[np.repeat(nodes, 8), np.tile(np.arange(8), 150)] = 90000 + 1000 * k
        key, elig = S.node_keys(cap, used, ndev, cls)
        states.append(S.filter_pass(reqs, key, elig))
    with gsx.Snapshot(sc.n_nodes, 8, K * 1200) as s:
        s.node_upsert(sc.gs_nodes())
        done = threading.Event()
        seen = set()

        def writer():
            for k, ev in enumerate(batches):
                (s.pods_apply if k % 2 else s.pods_apply_async)(ev)
            s.sync()
            done.set()

        def reader():
            last = 0
            while True:
                finished = done.is_set()
                bits = gsx.unpack_bitmap(s.filter_batch(reqs), sc.n_nodes)
                match = [j for j in range(last, K + 1) if (bits == states[j]).all()]
                assert match, "a filter saw a state that is no prefix of the event batches"
                last = match[0]                                   # and time does not run backwards
                seen.add(last)
                if finished:
                    assert last == K
                    return

        _run_threads([writer, reader, reader, reader])
        assert K in seen


def test_in_place_filters_beside_single_binds_see_a_prefix_of_the_binds():
    """The request-at-a-time routes under concurrency: one thread binds pods one by one (the one-warp kernel, in place in
    the page-locked block), three threads filter all the while -- the compact form into their own page-locked buffers
    (kernels writing host memory in place) and a small candidate list (in place in the context's block).  Every answer
    must be the answer for the state after some whole number of binds, and time must not run backwards."""
    sc = synth.make_cluster(64, 4, 16384, running=0, seed=5, quirks=False)
    cap, ndev, cls = sc.dev_capacity(), sc.dev_count(), sc.node_class()
    reqs = np.array([1, 5000, 9000, 16384], np.uint64)
    K = 120
    rng = np.random.default_rng(5)
    binds = synth.bind_records(rng.integers(0, 8, K), rng.choice([4096, 8192], K), first_slot=0)   # eight hot nodes
    used = np.zeros((sc.n_nodes, 4), np.uint64)
    table = np.zeros(K, synth.POD_DT)
    states = []
    key, elig = S.node_keys(cap, used, ndev, cls)
    states.append(S.filter_pass(reqs, key, elig))
    want_dev = []
    for k in range(K):
        want_dev.append(int(S.bind_sequential(binds[k:k + 1], cap, used, ndev, cls, table)[0]))
        key, elig = S.node_keys(cap, used, ndev, cls)
        states.append(S.filter_pass(reqs, key, elig))
    sub = np.arange(16, dtype=np.uint32)
    with gsx.Snapshot(sc.n_nodes, 4, K) as s:
        s.node_upsert(sc.gs_nodes())
        done = threading.Event()

        def writer():
            for k in range(K):
                assert int(s.bind_batch(binds[k:k + 1])[0]) == want_dev[k], k
            done.set()

        def reader(kind):
            def run():
                h_req = gsx.PinnedBuffer(len(reqs), np.uint64)
                h_rank = gsx.PinnedBuffer((1, sc.n_nodes), np.uint16)
                h_cut = gsx.PinnedBuffer(len(reqs), np.uint16)
                h_req.array[:] = reqs
                last = 0
                while True:
                    finished = done.is_set()
                    if kind == "compact":
                        rank, cut = s.filter_batch_compact(h_req.array, out_rank=h_rank.array, out_cutoff=h_cut.array)
                        bits = rank[0][None, :] >= cut[:, None]
                        match = [j for j in range(last, K + 1) if (bits == states[j]).all()]
                    else:
                        bits = gsx.unpack_bitmap(s.filter_batch_subset(reqs, sub), len(sub))
                        match = [j for j in range(last, K + 1) if (bits == states[j][:, :16]).all()]
                    assert match, "a filter saw a state that is no prefix of the binds"
                    last = match[0]
                    if finished:
                        assert (bits == (states[K] if kind == "compact" else states[K][:, :16])).all()
                        for b in (h_req, h_rank, h_cut):
                            b.free()
                        return
            return run

        _run_threads([writer, reader("compact"), reader("subset"), reader("compact")])
        assert (s.read_used() == used).all()


def test_two_snapshots_pipeline_on_one_gpu_from_two_threads():
    """Two snapshots, each driven from its own host thread on its own stream with overlapping filters switched
    on, share cuda:0: their kernels interleave on the SMs, and each keeps to its own work buffers and counters."""
    import torch

    scs = [synth.make_config("C3"), synth.make_cluster(9000, 8, 184320, running=60000, seed=77, quirks=True)]
    P, rounds = 3000, 40
    results = [None, None]
    with H.load_snapshot(scs[0]) as s0, H.load_snapshot(scs[1]) as s1:
        snaps = [s0, s1]
        start = threading.Barrier(2)

        def worker(t):
            def run():
                s, sc = snaps[t], scs[t]
                s.set_option(gsx.capi.OPT_PIPELINE_FILTERS, 1)
                stride = (s.bitmap_words + 7) // 8 * 8
                st = torch.cuda.Stream()
                batches = [synth.make_requests(P, "adversarial" if i % 2 else "catalogue", 900 + 10 * t + i).astype(np.uint64) for i in range(4)]
                d_reqs = [torch.from_numpy(b.view(np.int64)).cuda() for b in batches]
                outs = [torch.full((P, stride), -1, dtype=torch.int32, device="cuda") for _ in range(4)]
                torch.cuda.synchronize()
                start.wait()
                for i in range(rounds):
                    s.filter_batch_device(d_reqs[i % 4].data_ptr(), P, outs[i % 4].data_ptr(), stride, st.cuda_stream)
                st.synchronize()
                results[t] = (batches, [o.cpu().numpy().view(np.uint32) for o in outs], stride)
            return run

        _run_threads([worker(0), worker(1)])
    for t in range(2):
        cap, used, ndev, cls, _ = H.soa_state(scs[t])
        key, elig = S.node_keys(cap, used, ndev, cls)
        batches, outs, stride = results[t]
        for b, o in zip(batches, outs):
            assert (o == S.pack_bitmap(S.filter_pass(b, key, elig), stride)).all(), t


def test_pipelined_filters_beside_a_kernel_that_fills_the_gpu():
    """Overlapping filters while an unrelated stream keeps every SM busy with large element-wise kernels: the
    kernels of a batch may have to wait for SM slots, which delays them and changes nothing else."""
    import torch

    sc = synth.make_config("C3")
    cap, used, ndev, cls, _ = H.soa_state(sc)
    key, elig = S.node_keys(cap, used, ndev, cls)
    P = 5000
    with H.load_snapshot(sc) as s:
        s.set_option(gsx.capi.OPT_PIPELINE_FILTERS, 1)
        stride = (s.bitmap_words + 7) // 8 * 8
        st, side = torch.cuda.Stream(), torch.cuda.Stream()
        batches = [synth.make_requests(P, "adversarial", 40 + i).astype(np.uint64) for i in range(3)]
        d_reqs = [torch.from_numpy(b.view(np.int64)).cuda() for b in batches]
        outs = [torch.full((P, stride), -1, dtype=torch.int32, device="cuda") for _ in range(3)]
        big = torch.zeros(1 << 28, dtype=torch.float32, device="cuda")          # 1 GiB: ~0.3 ms per pass over it
        torch.cuda.synchronize()
        with torch.cuda.stream(side):
            for _ in range(60):
                big.mul_(1.0001).add_(1.0)
        for i in range(45):
            s.filter_batch_device(d_reqs[i % 3].data_ptr(), P, outs[i % 3].data_ptr(), stride, st.cuda_stream)
        st.synchronize()
        side.synchronize()
        for b, o in zip(batches, outs):
            assert (o.cpu().numpy().view(np.uint32) == S.pack_bitmap(S.filter_pass(b, key, elig), stride)).all()


def test_filter_on_another_stream_is_ordered_against_writers_without_host_syncs():
    """A device-buffer filter left running on the caller's stream, then pod events (which run on the snapshot's own
    stream), then another filter: the library orders the three on the GPU; the caller synchronises nothing."""
    import torch

    sc = synth.make_config("C3")
    cap, used, ndev, cls, _ = H.soa_state(sc)
    key0, elig = S.node_keys(cap, used, ndev, cls)
    P = 6000
    reqs = synth.make_requests(P, "adversarial", 3).astype(np.uint64)
    rng = np.random.default_rng(1)
    nodes = rng.choice(sc.n_nodes, 3000, replace=False)
    ev = synth.add_events(sc.n_pods + np.arange(3000), nodes, np.zeros(3000, np.int16), np.full(3000, 150000))
    with H.load_snapshot(sc, extra_pods=3000) as s:
        st = torch.cuda.Stream()
        stride = (s.bitmap_words + 7) // 8 * 8
        d_req = torch.from_numpy(reqs.view(np.int64)).cuda()
        outs = [torch.full((P, stride), -1, dtype=torch.int32, device="cuda") for _ in range(2)]
        torch.cuda.synchronize()
        for rep in range(3):
            if rep:
                s.pods_apply(synth.remove_events(sc.n_pods + np.arange(3000)))
            s.filter_batch_device(d_req.data_ptr(), P, outs[0].data_ptr(), stride, st.cuda_stream)
            (s.pods_apply if rep % 2 else s.pods_apply_async)(ev)        # no st.synchronize() in between
            s.filter_batch_device(d_req.data_ptr(), P, outs[1].data_ptr(), stride, st.cuda_stream)
            st.synchronize()
            used1 = s.read_used()
            key1, _ = S.node_keys(cap, used1, ndev, cls)
            assert (key1 != key0).any()
            assert (outs[0].cpu().numpy().view(np.uint32) == S.pack_bitmap(S.filter_pass(reqs, key0, elig), stride)).all(), rep
            assert (outs[1].cpu().numpy().view(np.uint32) == S.pack_bitmap(S.filter_pass(reqs, key1, elig), stride)).all(), rep


def test_bind_slot_rules():
    with gsx.Snapshot(4, 2, 16) as s:
        s.upsert_nodes([0, 1, 2], [100, 100, 100], [2, 2, 2], [0, 0, 0])
        s.pods_apply(synth.add_events([0], [0], [0], [60]))                       # slot 0 lives on node 0
        b = synth.bind_records([1, 2], [10, 10], first_slot=5)
        b["slot"] = [5, 5]
        with pytest.raises(gsx.GpushareError) as e:                               # one slot, two binds of the batch
            s.bind_batch(b)
        assert e.value.code == gsx.capi.GS_ERR_INVALID
        assert s.read_used().tolist() == [[60, 0], [0, 0], [0, 0]]                # nothing ran
        b = synth.bind_records([1, 0, 2], [10, 10, 10], first_slot=0)
        b["slot"] = [0, 1, 2]                                                     # slot 0 belongs to node 0's pod
        got = s.bind_batch(b)
        assert got.tolist() == [gsx.capi.BIND_SLOT_BUSY, 0, 0]
        assert s.read_used().tolist() == [[70, 0], [0, 0], [10, 0]]
        assert s.read_pods(0, 1)["mem"].tolist() == [60]                          # untouched
        # the same node: the entry is replaced (the same UID bound again), not counted twice
        b = synth.bind_records([0], [30], first_slot=0)
        assert s.bind_batch(b).tolist() == [0]                                    # 30 free on device 0 fits 30 exactly
        assert s.read_used().tolist()[0] == [40, 0]
        s.rebuild()
        assert s.read_used().tolist() == [[40, 0], [0, 0], [10, 0]]


def test_grid_sizing_uses_the_contexts_sm_count():
    import torch

    with gsx.Snapshot(8, 2, 8) as s:
        assert s.sm_count == torch.cuda.get_device_properties(0).multi_processor_count


def test_binds_to_different_nodes_overlap_in_the_apiserver_step():
    """Allocate holds the NODE's lock across Pods.Update and Pods.Bind (pkg/cache/nodeinfo.go:141-184): a bind that
    is inside its apiserver step must not delay a bind to another node, and must delay a bind to the same node."""
    with X.Extender(8, 2, 64) as xe:
        for n in ("A", "B"):
            xe.set_node(n, mem=200, count=2)
        for i, node in enumerate(["A", "B", "A"]):
            xe.list_pod(X.Pod(f"p{i}", uid=f"u{i}", limits=[60]))
        entered, release = threading.Event(), threading.Event()
        order = []

        def hook(key, node, mode):
            order.append(("api", key, node))
            if key == "default/p0":
                entered.set()
                assert release.wait(60)

        xe.set_api_hook(hook)
        res = {}

        def bind(i, node):
            def run():
                res[i] = xe.bind(f"p{i}", "default", f"u{i}", node)
                order.append(("done", i))
            return run

        t0 = threading.Thread(target=bind(0, "A"))
        t0.start()
        assert entered.wait(60)
        t1 = threading.Thread(target=bind(1, "B"))
        t1.start()
        t1.join(60)
        assert not t1.is_alive() and res[1] == (0, "")                   # B went through while A's bind is held
        t2 = threading.Thread(target=bind(2, "A"))
        t2.start()
        t2.join(0.5)
        assert t2.is_alive()                                             # the second bind to A waits for the first
        release.set()
        t0.join(60)
        t2.join(60)
        assert res[0] == (0, "") and res[2] == (1, "")                   # device 0 has 40 left after p0: p2 goes to device 1
        xe.set_api_hook(None)
        assert xe.inspect_node("A") == ([100, 100], [60, 60])
        assert order.index(("done", 1)) < order.index(("done", 0))


def test_a_drain_of_the_work_queue_is_one_device_batch():
    """syncPod handles one key at a time in the reference (controller.go:166-205); here every key of a drain, and
    every pod of a BuildCache, reaches the GPU in one gs_pods_apply_async call: two kernel launches, however many
    pods."""
    oc = O.Cluster()
    with X.Extender(64, 4, 4096) as xe:
        for i in range(40):
            xe.set_node(f"n{i}", mem=4 * 16384, count=4)
            oc.set_node(f"n{i}", mem=4 * 16384, count=4)
        rng = np.random.default_rng(0)
        pods = []
        for j in range(600):
            kw = dict(uid=f"u{j}", node=f"n{int(rng.integers(0, 40))}", phase=X.RUNNING,
                      annotations={O.IDX: int(rng.integers(0, 4)), O.POD: int(rng.choice([256, 1024, 2048]))}, limits=[1024])
            pods.append((X.Pod(f"p{j}", **kw), O.Pod(f"p{j}", **kw)))
        for px, po in pods[:300]:
            xe.list_pod(px)
            oc.list_pod(po)
        before = xe.launch_count
        assert xe.build_cache() is None and oc.build_cache() is None
        assert xe.launch_count - before == 2
        for px, po in pods[300:]:
            xe.list_pod(px)
            oc.list_pod(po)
            xe.informer_add(px)
            oc.informer_add(po)
        assert xe.queue_len == 300
        before = xe.launch_count
        assert xe.drain() is None and oc.drain() is None
        assert xe.launch_count - before == 2
        for i in range(40):
            assert xe.inspect_node(f"n{i}") == oc.inspect_node(f"n{i}")
