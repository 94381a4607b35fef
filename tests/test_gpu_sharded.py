"""Node-axis sharding on GPUs (BASELINE config 4's shape): the shards' kernels and the routing
against the unsharded restatement.

* on one GPU: G shard snapshots side by side on cuda:0, their results assembled in-process (no collective) --
  this runs wherever one GPU is visible;
* on >= 2 GPUs: two NCCL ranks, one process per GPU, the gather device to device (skipped when fewer GPUs are visible;
  tests/test_sharding_gloo.py covers the same collectives on CPU).
"""
import importlib
import os
import socket
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
if "gsx" not in sys.modules:  # spawned workers import this module without conftest
    sys.modules["gsx"] = importlib.import_module("gpushare-scheduler-extender_b200")

import numpy as np
import pytest
import torch

import gsx
from gsx import sharding, synth
from oracle import soa_reference as S

pytestmark = pytest.mark.gpu


def _scenario(sc, P, B):
    reqs = synth.make_requests(P, "adversarial", 5).astype(np.uint64)
    rng = np.random.default_rng(3)
    nodes = rng.integers(0, sc.n_nodes, B)
    hot = rng.integers(0, sc.n_nodes, max(1, sc.n_nodes // 100))
    pick = rng.random(B) < 0.5
    nodes[pick] = rng.choice(hot, int(pick.sum()))                # half the binds queue up on 1 % of the nodes
    binds = synth.bind_records(nodes, rng.choice(np.r_[synth.CATALOGUE[:8], [0]], B), first_slot=sc.n_pods)
    ev = np.concatenate([synth.add_events(sc.n_pods + B + np.arange(300), rng.integers(0, sc.n_nodes, 300),
                                          rng.integers(0, sc.max_devs, 300), rng.choice(synth.CATALOGUE[:6], 300)),
                         synth.remove_events(rng.integers(0, sc.n_pods, 500))])
    return reqs, binds, ev


def _globalise(sh, passes):
    """[P][N] bool -> shard-major packed rows: shard g's nodes start at word g * words_per_shard."""
    W = sh.world * sh.words_per_shard
    out = np.zeros((passes.shape[0], W * 32), dtype=bool)
    for g in range(sh.world):
        lo, hi = sh.node_range(g)
        out[:, g * sh.per_shard: g * sh.per_shard + (hi - lo)] = passes[:, lo:hi]
    return S.pack_bitmap(out, W)


def _unsharded(sc, reqs, binds, ev, extra):
    cap, ndev, cls = sc.dev_capacity(), sc.dev_count(), sc.node_class()
    recs = sc.gs_pod_records()
    recs["flags"][~S.legal(recs, ndev)] = 0
    used = S.used_from_records(recs, sc.n_nodes, sc.max_devs, ndev)
    table = np.zeros(sc.n_pods + extra, synth.POD_DT)
    table[: sc.n_pods] = recs
    key, elig = S.node_keys(cap, used, ndev, cls)
    p0 = S.filter_pass(reqs, key, elig)
    devs = S.bind_sequential(binds, cap, used, ndev, cls, table)
    key, elig = S.node_keys(cap, used, ndev, cls)
    p1 = S.filter_pass(reqs, key, elig)
    S.apply_events(ev, table, used, ndev)
    key, elig = S.node_keys(cap, used, ndev, cls)
    p2 = S.filter_pass(reqs, key, elig)
    return p0, devs, p1, p2, np.where(elig, np.searchsorted(np.sort(reqs), key, side="right"), 0)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_shards_on_one_gpu_match_unsharded(world):
    """C3 (heterogeneous capacities) split into `world` node ranges, one snapshot per range on cuda:0."""
    sc = synth.make_config("C3")
    P, B, extra = 1500, 1200, 4096
    reqs, binds, ev = _scenario(sc, P, B)
    p0, devs, p1, p2, rank_want = _unsharded(sc, reqs, binds, ev, extra)
    sh = sharding.NodeSharding(sc.n_nodes, world)
    snaps = [sharding.load_shard(sh, g, sc, extra_pods=extra) for g in range(world)]
    try:
        engines = [sharding.SnapshotEngine(s) for s in snaps]
        d_req = torch.from_numpy(reqs.view(np.int64)).cuda()

        def gathered():
            parts = [e.filter(d_req, sh.stride_words).clone() for e in engines]
            torch.cuda.synchronize()
            return np.stack([p.cpu().numpy().view(np.uint32) for p in parts])

        assert (sh.assemble_bitmap(gathered()) == _globalise(sh, p0)).all()
        got = np.full(B, np.iinfo(np.int32).min, np.int32)
        for g in range(world):
            mine, pos = sh.route(binds, g)
            if len(mine):
                got[pos] = engines[g].bind(mine)
        assert (got == devs).all()
        assert (sh.assemble_bitmap(gathered()) == _globalise(sh, p1)).all()
        for g in range(world):
            mine, _ = sh.route_events(ev, g)
            engines[g].apply_events(mine)
        assert (sh.assemble_bitmap(gathered()) == _globalise(sh, p2)).all()
        ranks, cuts = [], []
        for e in engines:
            r, c = e.filter_compact(d_req, sh.per_shard)
            torch.cuda.synchronize()
            ranks.append(r.cpu().numpy().view(np.uint16).copy())
            cuts.append(c.cpu().numpy().view(np.uint16).copy())
        for c in cuts[1:]:
            assert (c == cuts[0]).all()                               # cutoffs depend on the requests only
        rank = sh.assemble_rank(np.stack(ranks))
        assert (rank == rank_want).all()
        assert ((rank[None, :] >= cuts[0][:, None]) == p2).all()
    finally:
        for s in snaps:
            s.close()


def _nccl_worker(rank, world, port, out_dir):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        sc = synth.make_config("C3")
        P, B, extra = 1500, 1200, 4096
        reqs, binds, ev = _scenario(sc, P, B)
        sh = sharding.NodeSharding(sc.n_nodes, world)
        snap = sharding.load_shard(sh, rank, sc, extra_pods=extra, device=rank)
        cl = sharding.ShardedCluster(sh, rank, sharding.SnapshotEngine(snap), dist, device=torch.device("cuda", rank))
        req_t = cl.broadcast_requests(reqs if rank == 0 else None)
        g0 = cl.filter_device(req_t).cpu().numpy().view(np.uint32).copy()
        devs = cl.bind(binds if rank == 0 else None)
        g1 = cl.filter_device(req_t).cpu().numpy().view(np.uint32).copy()
        cl.apply_events(ev if rank == 0 else None)
        g2 = cl.filter_device(req_t).cpu().numpy().view(np.uint32).copy()
        rk, cut = cl.filter_compact_device(req_t)
        rk, cut = rk.cpu().numpy().view(np.uint16).copy(), cut.cpu().numpy().view(np.uint16).copy()
        # the same with the gather done by the rank kernel itself (peer stores into symmetric memory), several calls
        # in a row so that both of its alternating buffers are used and reused
        cl.enable_peer_gather()
        peer_ok = True
        for _ in range(5):
            rk2, cut2 = cl.filter_compact_device(req_t)
            peer_ok &= bool((rk2.cpu().numpy().view(np.uint16) == rk).all() and (cut2.cpu().numpy().view(np.uint16) == cut).all())
        np.savez(os.path.join(out_dir, f"r{rank}.npz"), g0=g0, devs=devs, g1=g1, g2=g2, rk=rk, cut=cut, peer_ok=np.array(peer_ok))
        snap.close()
    finally:
        dist.destroy_process_group()


def test_two_nccl_ranks_match_unsharded(tmp_path):
    """C3 over two GPUs: requests broadcast, per-shard kernels, all-gather of the shard bitmaps device to device,
    binds to the owner rank only, events routed; every rank ends with the unsharded restatement's answers."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp

    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_nccl_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    sc = synth.make_config("C3")
    P, B, extra = 1500, 1200, 4096
    reqs, binds, ev = _scenario(sc, P, B)
    p0, devs, p1, p2, rank_want = _unsharded(sc, reqs, binds, ev, extra)
    sh = sharding.NodeSharding(sc.n_nodes, world)
    res = [np.load(tmp_path / f"r{r}.npz") for r in range(world)]
    for k in ("g0", "devs", "g1", "g2", "rk", "cut"):
        assert (res[0][k] == res[1][k]).all(), k
    r0 = res[0]
    assert (sh.assemble_bitmap(r0["g0"]) == _globalise(sh, p0)).all()
    assert (r0["devs"] == devs).all()
    assert (sh.assemble_bitmap(r0["g1"]) == _globalise(sh, p1)).all()
    assert (sh.assemble_bitmap(r0["g2"]) == _globalise(sh, p2)).all()
    rank = sh.assemble_rank(r0["rk"])
    assert (rank == rank_want).all() and ((rank[None, :] >= r0["cut"][:, None]) == p2).all()
    assert all(bool(r["peer_ok"]) for r in res), "the kernel-side gather over peer memory differs from the NCCL gather"
