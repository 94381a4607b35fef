"""gs_pods_apply_static -- pod events for hosts that are not allowed to allocate GPU memory.

Some embedders of this library hand the whole GPU to one memory owner at start-up and forbid cudaMalloc
afterwards (so a scheduling call can never fail or stall inside the allocator), or run where the implicit
device-wide synchronisation of cudaMalloc / cudaFree cannot be tolerated.  For them the library offers one
entry point that allocates nothing: the caller supplies every buffer as a device address and the kernel's
scratch table is a fixed array inside the library.

This test drives that entry point through the C ABI, twice in a row on the same buffers (the fixed array
keeps what the previous call left, so the second call proves it is reset), and requires the pod table and
used[] to be bit-identical to what the ordinary, allocating gs_pods_apply produces from the same events.
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
        for ev in batches:                  # two calls in a row
            ref.pods_apply(ev)              # the existing entry point (stages and may allocate)
            d_ev = torch.from_numpy(ev.view(np.uint8)).cuda()   # the caller's own device copy of the events
            torch.cuda.synchronize()        # ... complete before the call's stream reads it
            capi.pods_apply_static(0, d_ev.data_ptr(), len(ev), ptr["pods"], n_slots, ptr["nodeinfo"], n_nodes,
                                   n_devs, ptr["used"], stream.cuda_stream)
            stream.synchronize()
            assert (mine.read_used() == ref.read_used()).all()
            assert (mine.read_pods(0, n_slots) == ref.read_pods(0, n_slots)).all()
        # and the table it left adds up to the used[] it left (numpy restatement of the rebuild)
        table = mine.read_pods(0, n_slots)
        assert (S.used_from_records(table, n_nodes, n_devs, sc.dev_count()) == mine.read_used()).all()


def test_static_scratch_entry_point_refuses_oversized_tables():
    with gsx.Snapshot(4, 2, 4) as s:
        s.upsert_nodes([0], [100], [2], [0])
        ptr = s.device_ptrs()
        with pytest.raises(gsx.GpushareError) as e:
            capi.pods_apply_static(0, ptr["pods"], 1, ptr["pods"], capi.STATIC_MAX_SLOTS + 1, ptr["nodeinfo"], 4, 2,
                                   ptr["used"], 0)
        assert e.value.code == capi.GS_ERR_CAPACITY and "at most" in str(e.value)
