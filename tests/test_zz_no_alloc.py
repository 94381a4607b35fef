"""gs_pods_apply_static -- pod events for hosts that are not allowed to allocate GPU memory.

Some embedders of this library hand the whole GPU to one memory owner at start-up and forbid cudaMalloc
afterwards (so a scheduling call can never fail or stall inside the allocator), or run where the implicit
device-wide synchronisation of cudaMalloc / cudaFree cannot be tolerated.  For them the library offers one
entry point that allocates nothing: the caller supplies every buffer as a device address and the kernel's
scratch table is a fixed array inside the library.

This test drives that entry point through the C ABI, twice in a row on the same buffers (the fixed array
keeps what the previous call left, so the second call proves it is reset), and requires the pod table and
used[] to be bit-identical both to what the ordinary, allocating gs_pods_apply produces from the same events
and to the CPU restatement of the events applied one by one (the two entry points share their device code,
so agreeing with each other alone would not catch a mistake common to both).
It is collected last (file name) so that a failure here cannot hide any other test's result, and it builds
all of its own state, so it gives the same verdict alone in a fresh process and after the rest of the suite.
"""
import numpy as np
import pytest

import gsx
from gsx import capi, synth
from oracle import soa_reference as S

pytestmark = pytest.mark.gpu


def _batch(rng, n_events, n_slots, n_nodes, n_devs):
    ev = np.zeros(n_events, synth.EVENT_DT)
    ev["slot"] = rng.integers(0, n_slots, n_events)          # repeated slots inside one batch on purpose
    ev["op"] = rng.random(n_events) < 0.35
    ev["node_idx"] = rng.integers(0, n_nodes + 1, n_events)  # a few beyond the node table
    ev["dev"] = rng.integers(-1, n_devs + 1, n_events)       # a few illegal device ids
    ev["mem"] = rng.choice(synth.CATALOGUE, n_events)
    ev["counted"] = rng.random(n_events) < 0.9
    return ev


def test_static_scratch_entry_point_matches_allocating_one():
    import torch

    rng = np.random.default_rng(20261106)
    n_nodes, n_devs, n_slots = 300, 8, 5000
    sc = synth.make_cluster(n_nodes, n_devs, 184320, 0, seed=77, quirks=True)
    batches = [_batch(rng, 4000, n_slots, n_nodes, n_devs), _batch(rng, 2500, n_slots, n_nodes, n_devs)]

    with gsx.Snapshot(n_nodes, n_devs, n_slots) as ref, gsx.Snapshot(n_nodes, n_devs, n_slots) as mine:
        ref.node_upsert(sc.gs_nodes())
        mine.node_upsert(sc.gs_nodes())
        ptr = mine.device_ptrs()            # buffers the "caller" already owns: the second snapshot's arrays
        stream = torch.cuda.Stream()
        cpu_table = np.zeros(n_slots, synth.POD_DT)
        cpu_used = np.zeros((n_nodes, n_devs), np.uint64)
        for ev in batches:                  # two calls in a row
            S.apply_events(ev, cpu_table, cpu_used, sc.dev_count())   # the restatement, event by event
            ref.pods_apply(ev)              # the existing entry point (stages and may allocate)
            d_ev = torch.from_numpy(ev.view(np.uint8)).cuda()   # the caller's own device copy of the events
            torch.cuda.synchronize()        # ... complete before the call's stream reads it
            capi.pods_apply_static(0, d_ev.data_ptr(), len(ev), ptr["pods"], n_slots, ptr["nodeinfo"], n_nodes,
                                   n_devs, ptr["used"], stream.cuda_stream)
            stream.synchronize()
            assert (mine.read_used() == ref.read_used()).all()
            assert (mine.read_pods(0, n_slots) == ref.read_pods(0, n_slots)).all()
            assert (mine.read_used() == cpu_used).all()
            assert (mine.read_pods(0, n_slots) == cpu_table).all()
        # and the table it left adds up to the used[] it left (numpy restatement of the rebuild)
        table = mine.read_pods(0, n_slots)
        assert (S.used_from_records(table, n_nodes, n_devs, sc.dev_count()) == mine.read_used()).all()


def test_static_events_on_snapshot_arrays_survive_rebuild_and_order_with_filters():
    """The arrays of a snapshot are legitimate caller-owned buffers for this entry point.  Two things must hold
    then: a later gs_rebuild counts the records it wrote (also in slots no other call ever touched), and with
    overlapping filters switched on, a filter queued after the call (with gs_snapshot_external_write in between)
    sees what it wrote."""
    import torch

    rng = np.random.default_rng(5)
    n_nodes, n_devs, n_slots = 640, 8, 9000
    sc = synth.make_cluster(n_nodes, n_devs, 184320, 0, seed=78, quirks=False)
    reqs = rng.integers(1, 184321, 2000).astype(np.uint64)
    with gsx.Snapshot(n_nodes, n_devs, n_slots) as s:
        s.node_upsert(sc.gs_nodes())
        s.set_option(capi.OPT_PIPELINE_FILTERS, 1)
        ptr = s.device_ptrs()
        st = torch.cuda.Stream()
        stride = (s.bitmap_words + 7) // 8 * 8
        d_req = torch.from_numpy(reqs.view(np.int64)).cuda()
        outs = [torch.zeros((len(reqs), stride), dtype=torch.int32, device="cuda") for _ in range(3)]
        # fill most of every device of half of the nodes, in the highest slots of the table only
        nodes = np.repeat(np.arange(0, n_nodes, 2), n_devs)
        devs = np.tile(np.arange(n_devs), n_nodes // 2)
        ev = synth.add_events(n_slots - 1 - np.arange(len(nodes)), nodes, devs, np.full(len(nodes), 150000))
        d_ev = torch.from_numpy(ev.view(np.uint8)).cuda()
        torch.cuda.synchronize()
        s.filter_batch_device(d_req.data_ptr(), len(reqs), outs[0].data_ptr(), stride, st.cuda_stream)
        s.filter_batch_device(d_req.data_ptr(), len(reqs), outs[1].data_ptr(), stride, st.cuda_stream)
        capi.pods_apply_static(0, d_ev.data_ptr(), len(ev), ptr["pods"], n_slots, ptr["nodeinfo"], n_nodes, n_devs,
                               ptr["used"], st.cuda_stream)
        s.external_write(st.cuda_stream)
        s.filter_batch_device(d_req.data_ptr(), len(reqs), outs[2].data_ptr(), stride, st.cuda_stream)
        st.synchronize()
        cap, ndev, cls = sc.dev_capacity(), sc.dev_count(), sc.node_class()
        used = np.zeros((n_nodes, n_devs), np.uint64)
        key0, elig = S.node_keys(cap, used, ndev, cls)
        used[nodes, devs] = 150000
        key1, _ = S.node_keys(cap, used, ndev, cls)
        before = S.pack_bitmap(S.filter_pass(reqs, key0, elig), stride)
        after = S.pack_bitmap(S.filter_pass(reqs, key1, elig), stride)
        assert (before != after).any()
        assert (outs[0].cpu().numpy().view(np.uint32) == before).all()
        assert (outs[1].cpu().numpy().view(np.uint32) == before).all()
        assert (outs[2].cpu().numpy().view(np.uint32) == after).all()
        assert (s.read_used() == used).all()
        s.rebuild()                              # reads the whole table, the high slots included
        assert (s.read_used() == used).all()


def test_static_scratch_entry_point_skips_what_the_table_cannot_hold():
    """A caller-built node table may claim more devices than used[] has columns, and an event may carry an op code
    that is neither add nor remove: both are skipped, nothing is written outside the arrays."""
    import torch

    n_nodes, n_devs, n_slots = 4, 2, 16
    nodeinfo = torch.tensor([6, 2, 2, 2], dtype=torch.int32, device="cuda")       # node 0 claims 6 devices, stride is 2
    pods = torch.zeros(n_slots * 16, dtype=torch.uint8, device="cuda")
    used = torch.zeros(n_nodes * n_devs + 8, dtype=torch.int64, device="cuda")     # 8 guard words behind the array
    ev = synth.add_events([0, 1, 2, 3], [0, 0, 1, 1], [1, 5, 0, 1], [10, 20, 30, 40])
    ev["op"][3] = 7
    d_ev = torch.from_numpy(ev.view(np.uint8)).cuda()
    torch.cuda.synchronize()
    capi.pods_apply_static(0, d_ev.data_ptr(), len(ev), pods.data_ptr(), n_slots, nodeinfo.data_ptr(), n_nodes, n_devs,
                           used.data_ptr(), 0)
    torch.cuda.synchronize()
    assert used.cpu().tolist() == [0, 10, 30, 0, 0, 0, 0, 0] + [0] * 8
    table = pods.cpu().numpy().view(synth.POD_DT)
    assert table["flags"].tolist()[:4] == [3, 0, 3, 0]


def test_static_scratch_entry_point_refuses_oversized_tables():
    with gsx.Snapshot(4, 2, 4) as s:
        s.upsert_nodes([0], [100], [2], [0])
        ptr = s.device_ptrs()
        with pytest.raises(gsx.GpushareError) as e:
            capi.pods_apply_static(0, ptr["pods"], 1, ptr["pods"], capi.STATIC_MAX_SLOTS + 1, ptr["nodeinfo"], 4, 2,
                                   ptr["used"], 0)
        assert e.value.code == capi.GS_ERR_CAPACITY and "at most" in str(e.value)
